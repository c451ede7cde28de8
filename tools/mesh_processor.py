#!/usr/bin/env python
"""The reference's offline mesh tool (tool/meshProcessor.cpp) on top of librtrt.so: read an .obj or .bin, order the triangles by the
60-bit Morton code of their centroid ON THE DEVICE, write `uint32 n; Triangle[n]`.  Needs a GPU (the sort is the product's
bvh_presort.cu path).

    python tools/mesh_processor.py in.obj out.bin [--no-sort] [--legacy64]
"""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("input"); ap.add_argument("output")
    ap.add_argument("--no-sort", action="store_true", help="keep the input triangle order")
    ap.add_argument("--legacy64", action="store_true", help="64-byte triangles (positions only)")
    a = ap.parse_args()
    pkg = importlib.import_module("real-time-ray-tracing_b200")
    with pkg.RayTracer(64, 64) as rt:
        if a.input.lower().endswith(".obj"):
            rt.set_mesh_obj(a.input)
        else:
            rt.set_mesh_bin(a.input)
        rt.write_mesh_bin(a.output, presort=not a.no_sort, legacy64=a.legacy64)
    print(f"wrote {a.output} ({os.path.getsize(a.output)} bytes)")


if __name__ == "__main__":
    main()
