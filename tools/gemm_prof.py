"""One NT (bf16 and bf16x3) and one TN launch of the tensor-core contraction at LSTM-gate size, for ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tools.gemm_check import nt, tn
M, N, K = 16384, 1792, 2048
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5
C = torch.empty(M, N, device="cuda")
for prec in (2, 1):
    nt(prec, A, W, False, C)
P = torch.randn(M, N, device="cuda"); G = torch.zeros(N, K, device="cuda")
for prec in (2, 1):
    tn(prec, P, A, G)
torch.cuda.synchronize()
print("ok")
