"""Error and throughput of the stand-alone contraction entry points for LSTM-sized problems (run on the B200 box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from paths_b200 import _lib as L

lib = L.lib()
st = lambda: torch.cuda.current_stream().cuda_stream


def nt(prec, A, W, b_kn, C, bias=None, flags=0):
    M, K = A.shape
    N = W.shape[1] if b_kn else W.shape[0]
    L.check(lib.paths_gemm_nt(prec, A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0), int(b_kn), C.data_ptr(), C.stride(0),
                              bias.data_ptr() if bias is not None else None, M, N, K, flags, st()), "nt")


def tn(prec, P, Q, C):
    M, NI = P.shape
    NJ = Q.shape[1]
    L.check(lib.paths_gemm_tn(prec, P.data_ptr(), NI, Q.data_ptr(), NJ, C.data_ptr(), NJ, M, NI, NJ, st()), "tn")


def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    for (M, N, K) in [(16384, 1792, 2048), (16384, 1024, 256), (16384, 256, 1024), (16384, 384, 128)]:
        A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5
        ref = A.double() @ W.double().t()
        for prec in (0, 1, 2):
            C = torch.empty(M, N, device="cuda")
            nt(prec, A, W, False, C)
            err = float((C.double() - ref).abs().max() / ref.abs().max())
            ms = timeit(lambda: nt(prec, A, W, False, C))
            print(f"NT  M={M} N={N} K={K} prec={prec}: rel err {err:.2e}  {ms:.3f} ms  {2*M*N*K/ms/1e9:.1f} TFLOP/s (incl. pack+sync)")
        Wt = W.t().contiguous()
        for prec in (0, 1, 2):
            C = torch.empty(M, N, device="cuda")
            nt(prec, A, Wt, True, C)
            err = float((C.double() - ref).abs().max() / ref.abs().max())
            print(f"NN  M={M} N={N} K={K} prec={prec}: rel err {err:.2e}")
    for (M, NI, NJ) in [(16384, 1792, 2048), (16384, 1024, 256)]:
        P = torch.randn(M, NI, device="cuda"); Q = torch.randn(M, NJ, device="cuda")
        ref = P.double().t() @ Q.double()
        for prec in (0, 1, 2):
            C = torch.zeros(NI, NJ, device="cuda")
            tn(prec, P, Q, C)
            err = float((C.double() - ref).abs().max() / ref.abs().max())
            ms = timeit(lambda: tn(prec, P, Q, C))
            print(f"TN  M={M} NI={NI} NJ={NJ} prec={prec}: rel err {err:.2e}  {ms:.3f} ms  {2*M*NI*NJ/ms/1e9:.1f} TFLOP/s")
    A = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16); B = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
    ms = timeit(lambda: A @ B)
    print(f"cuBLAS bf16 8192^3: {ms:.3f} ms {2*8192**3/ms/1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    main()
