import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels, synth
M,K=7,896
x = synth.activations(M, K, torch.bfloat16, seed=M + K)
x[0,:8]=0; x[2]=0
xq, xs = kernels.per_token_quant_int8(x.cuda())
rq, rs = qo.per_token_quant_int8(x)
d = (xq.cpu() != rq).nonzero()
print("mismatches", d.shape[0])
for (r,c) in d[:10].tolist():
    xf = x[r,c].float().item(); am = x[r].float().abs().max().item()
    print(r,c,"x",xf,"absmax",am,"gpu",xq[r,c].item(),"ref",rq[r,c].item(),"prod",xf*(127/am), "scale gpu/ref", xs[r].item(), rs[r].item())
