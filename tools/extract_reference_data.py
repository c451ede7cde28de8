#!/usr/bin/env python
"""Extract the DATA tables the hot path needs from the reference tree into small binary files.

These are published data sets, not code:
  * Heitz et al. 2019 blue-noise sampler tables, 4-spp variant (src/blueNoiseRandGenData.h:15-39):
      sobol_256spp_256d u8[65536], scramblingTile u8[131072], rankingTile u8[131072]
  * Hosek-Wilkie sky model coefficient sets (src/skyData.h): skyDataSets f32[540], skyDataSetsRad f32[60],
      solarDatasets f32[1800], limbDarkeningDatasets f32[60]
Run in the build container (needs /root/reference):  python tools/extract_reference_data.py
Outputs: real-time-ray-tracing_b200/data/{blue_noise_4spp.bin,hosek_sky.bin}
"""
import os
import re
import sys

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "real-time-ray-tracing_b200", "data")


def arrays_in(text):
    out = {}
    for m in re.finditer(r"([\w:]+)\s*\[[^\]]*\]\s*=\s*\{(.*?)\};", text, flags=re.S):
        body = re.sub(r"//.*", "", m.group(2))
        out[m.group(1)] = [t for t in re.split(r"[,\s]+", body.strip()) if t]
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = open(os.path.join(REF, "src", "blueNoiseRandGenData.h")).read()
    sect = txt[txt.index("#if OPTIMIZED_BLUE_NOISE_SPP == 4"):]
    sect = sect[:sect.index("#endif")]
    arrs = arrays_in(sect)
    names = {k.split("::")[-1]: v for k, v in arrs.items()}
    sob = np.array([int(x) for x in names["h_sobol_256spp_256d"]], np.uint8)
    scr = np.array([int(x) for x in names["h_scramblingTile"]], np.uint8)
    rnk = np.array([int(x) for x in names["h_rankingTile"]], np.uint8)
    assert sob.size == 65536 and scr.size == 131072 and rnk.size == 131072, (sob.size, scr.size, rnk.size)
    with open(os.path.join(OUT, "blue_noise_4spp.bin"), "wb") as f:
        f.write(sob.tobytes()); f.write(scr.tobytes()); f.write(rnk.tobytes())

    sky = arrays_in(open(os.path.join(REF, "src", "skyData.h")).read())
    def fl(name, n):
        a = np.array([float(x.rstrip("f")) for x in sky[name]], np.float32)
        assert a.size == n, (name, a.size)
        return a
    with open(os.path.join(OUT, "hosek_sky.bin"), "wb") as f:
        for name, n in [("skyDataSets", 540), ("skyDataSetsRad", 60), ("h_solarDatasets", 1800), ("h_limbDarkeningDatasets", 60)]:
            f.write(fl(name, n).tobytes())
    print("wrote", os.listdir(OUT))


if __name__ == "__main__":
    main()
