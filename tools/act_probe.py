"""Which float operation order do torch's CUDA normalize / sigmoid kernels use?  Bitwise comparison with candidates."""
import torch, torch.nn.functional as F
torch.manual_seed(0)
x = (torch.randn(2000000, 4, device="cuda") * torch.rand(2000000, 1, device="cuda") * 3)
ref_n = torch.linalg.vector_norm(x, dim=1)
ref_q = F.normalize(x)
sq = x * x
f64 = x.double()
def fma(a, b, c):  # float32 fma emulated in float64
    return (a.double() * b.double() + c.double()).float()
cands = {
    "((0+1)+2)+3": torch.sqrt(((sq[:, 0] + sq[:, 1]) + sq[:, 2]) + sq[:, 3]),
    "(0+1)+(2+3)": torch.sqrt((sq[:, 0] + sq[:, 1]) + (sq[:, 2] + sq[:, 3])),
    "fma chain 0,1,2,3": torch.sqrt(fma(x[:, 3], x[:, 3], fma(x[:, 2], x[:, 2], fma(x[:, 1], x[:, 1], sq[:, 0])))),
    "fma chain 3,2,1,0": torch.sqrt(fma(x[:, 0], x[:, 0], fma(x[:, 1], x[:, 1], fma(x[:, 2], x[:, 2], sq[:, 3])))),
    "(0+2)+(1+3)": torch.sqrt((sq[:, 0] + sq[:, 2]) + (sq[:, 1] + sq[:, 3])),
    "fma pairs (0,1)+(2,3)": torch.sqrt(fma(x[:, 1], x[:, 1], sq[:, 0]) + fma(x[:, 3], x[:, 3], sq[:, 2])),
    "double": f64.pow(2).sum(1).sqrt().float(),
}
for k, v in cands.items():
    print("norm %-24s mismatches %d" % (k, int((v != ref_n).sum())))
best = min(cands, key=lambda k: int((cands[k] != ref_n).sum()))
n = cands[best].clamp_min(1e-12)
print("normalize = x / norm(%s): mismatches %d; x * (1/norm): %d" % (best, int(((x / n[:, None]) != ref_q).sum()),
      int(((x * (1.0 / n)[:, None]) != ref_q).sum())))
print("normalize vs x / ref_norm:", int(((x / ref_n.clamp_min(1e-12)[:, None]) != ref_q).sum()))
y = torch.randn(4000000, device="cuda") * 6
rs = torch.sigmoid(y)
c = {"1/(1+exp(-x))": 1.0 / (1.0 + torch.exp(-y)), "exp(x)/(1+exp(x))": torch.exp(y) / (1 + torch.exp(y)),
     "double": torch.sigmoid(y.double()).float()}
for k, v in c.items():
    print("sigmoid %-20s mismatches %d" % (k, int((v != rs).sum())))
e = torch.exp(y)
print("exp vs correctly rounded:", int((e != torch.exp(y.double()).float()).sum()), "of", y.numel())
