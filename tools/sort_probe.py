"""Sort probe: the core's radix sort (rt_sort_pairs, device time of the sort kernels) against the call the reference makes
(cub::DeviceRadixSort::SortPairs over 32 bits, through oracle/_ref/libref_shim.so) on the same (key, index) pairs.

A digit pass moves 16 B per pair (8 B read, 8 B written) and the histogram kernel reads 4 B per pair once, so a sort over
P passes has 4 + 16 P algorithmic bytes per pair; GB/s below is that figure over the measured time.
"""
import ctypes as C
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import numpy as np
import cuda_raytracer_b200 as rt
import pyoracle as po

sizes = [int(float(v)) for v in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["1e6", "16e6", "64e6"])]
r = rt.Renderer(0)
shim = po.shim()
for n in sizes:
    rs = np.random.RandomState(n % 9973)
    base = rs.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
    cases = (("uniform32", base, (32,)), ("uniform15", base & np.uint32(0x7FFF), (15, 16)), ("runs15", np.sort(base & np.uint32(0x7FFF))[::-1].copy(), (15,)))
    if os.environ.get("PROBE_NO_CUB") == "1":
        cases = cases[:1]
    for label, keys, bit_list in cases:
        vals = np.arange(n, dtype=np.uint32)
        want = None
        for bits in bit_list:
            best = 1e30
            for rep in range(3):
                k, v = r.sort_pairs(keys, vals, bits)
                best = min(best, r.sort_device_ms)
            passes = (bits + 7) // 8
            gbs = n * (4 + 16 * passes) / best * 1e-6
            print(f"n {n:>10d} {label:10s} core  {bits:2d} bits ({passes} passes): {best:8.3f} ms  {n/best*1e-6:8.2f} Gpairs/s  {gbs:7.0f} GB/s algorithmic")
            order = np.argsort(keys, kind="stable") if want is None else want
            want = order
            assert (v == vals[order]).all() and (k == keys[order]).all(), "core sort wrong"
        if shim is not None and os.environ.get("PROBE_NO_CUB") != "1":
            k2, v2 = keys.copy(), vals.copy()
            ms = C.c_float()
            rc = shim.ref_cub_sort_pairs(k2, v2, n, 3, C.byref(ms))
            assert rc == 0, rc
            assert (v2 == vals[want]).all()
            print(f"n {n:>10d} {label:10s} cub   32 bits (reference call): {ms.value:8.3f} ms  {n/ms.value*1e-6:8.2f} Gpairs/s")
r.close()
