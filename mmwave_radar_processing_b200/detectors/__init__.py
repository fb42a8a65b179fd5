from .base import BaseCFAR1D, BaseCFAR2D
from .ca_cfar import CaCFAR1D, CaCFAR2D
from .go_so_cfar import GoCFAR1D, SoCFAR1D
from .os_cfar import OsCFAR1D, OsCFAR2D
