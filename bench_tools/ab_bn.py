"""A/B of BatchNorm entry points whose C signature did not change between two builds of the kernel library.
   python bench_tools/ab_bn.py libA.so libB.so"""
import ctypes
import sys

import torch

P = ctypes.c_void_p


def timeit(fn, iters=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    torch.cuda.init()
    libs = [ctypes.CDLL(p) for p in sys.argv[1:3]]
    st = P(torch.cuda.current_stream().cuda_stream)
    for rows, C in ((1254400, 256), (313600, 512), (78400, 1024), (1254400, 64)):
        x = torch.randn(rows, C, device="cuda").half()
        dz = torch.randn(rows, C, device="cuda").half()
        dx, dres = torch.empty_like(x), torch.empty_like(x)
        bits = torch.randint(0, 256, (rows * C // 8,), device="cuda", dtype=torch.uint8)
        mean, rstd, gamma = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda"), torch.ones(C, device="cuda")
        ds = torch.zeros(2 * C, device="cuda")
        res = []
        for L in libs:
            f = L.pdb_bn_bwd_apply
            f.argtypes = [P, P, P, P, P, P, ctypes.c_int, ctypes.c_int64, ctypes.c_int, P, P, P, P, ctypes.c_double, P, P, P]
            for with_dres in (True, False):
                t = timeit(lambda: f(dz.data_ptr(), None, bits.data_ptr(), x.data_ptr(), dx.data_ptr(), dres.data_ptr() if with_dres else None,
                                     1, rows, C, mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(), ds.data_ptr(), float(rows), None, None, st))
                res.append(round(t, 1))
        print((rows, C), "A +dres / A / B +dres / B us:", res)


if __name__ == "__main__":
    main()
