"""Regenerate the minimal TI .cfg files under configs/ from the reference snapshot.

Only the six commands that ConfigManager reads (channelCfg, adcCfg, adcbufCfg,
profileCfg, chirpCfg, frameCfg) are radar parameters; every other line of a TI
visualizer export is ignored by the parser (reference cfgManager.py:246-264), so
the files written here carry just those numbers plus a '%' header (lines that
contain '%' are skipped by the parser, reference cfgManager.py:234).

Run in the build container only (needs /root/reference):
    python tools/extract_cfgs.py
"""
import glob
import os
import sys

KEEP = ("channelCfg", "adcCfg", "adcbufCfg", "profileCfg", "chirpCfg", "frameCfg")


def main(ref_dir="/root/reference/configs", out_dir=None):
    out_dir = out_dir or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs")
    os.makedirs(out_dir, exist_ok=True)
    for path in sorted(glob.glob(os.path.join(ref_dir, "*.cfg"))):
        name = os.path.basename(path)
        rows = []
        with open(path) as fh:
            for line in fh:
                if "%" in line:
                    continue
                tok = line.split()
                if tok and tok[0] in KEEP:
                    rows.append(" ".join(tok))
        with open(os.path.join(out_dir, name), "w") as fh:
            fh.write("%% radar parameters of %s (parser-relevant commands only)\n" % name)
            fh.write("% consumed by mmwave_radar_processing_b200.config_managers.cfgManager\n")
            for r in rows:
                fh.write(r + "\n")
        print(name, len(rows))


if __name__ == "__main__":
    main(*sys.argv[1:])
