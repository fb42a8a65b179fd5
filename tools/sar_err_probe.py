import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
from test_gpu_sar_tc import sar_case, sar_direct, engine_with_mode
from conftest import rel_l2
for n in (125, 500, 2000, 8000):
    x, pos, dirs, lam = sar_case(90, n, 256, seed=n)
    want = sar_direct(x, pos, dirs, lam, np.arange(256))
    xs, ps, ds = torch.from_numpy(x).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(dirs).cuda()
    for mode in (1, 2):
        out = engine_with_mode(mode, 90).sar_image(xs, ps, ds, lam).cpu().numpy()
        print(n, mode, rel_l2(out, want), flush=True)
