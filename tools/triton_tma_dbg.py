"""debugging aid only (not product code): does a user-level TMA tensor load work on this box?  Dumps Triton's PTX."""
import torch, triton, triton.language as tl
from triton.tools.tensor_descriptor import TensorDescriptor

@triton.jit
def k(desc, out_ptr, BM: tl.constexpr, BN: tl.constexpr):
    t = desc.load([5, 32])
    offs = tl.arange(0, BM)[:, None] * BN + tl.arange(0, BN)[None, :]
    tl.store(out_ptr + offs, t)

x = (torch.arange(256 * 4096, device="cuda", dtype=torch.int32) % 30000).to(torch.int16).reshape(256, 4096)
out = torch.zeros(16 * 16, device="cuda", dtype=torch.int16)
d = TensorDescriptor.from_tensor(x, [16, 16])
h = k[(1,)](d, out, 16, 16)
torch.cuda.synchronize()
print("triton TMA ok:", out[:4].tolist(), x[5, 32:36].tolist())
ptx = h.asm["ptx"]
import re
for l in ptx.splitlines():
    if "tensor" in l or "mbarrier" in l or "fence" in l or ".target" in l or ".version" in l or "%rd1" in l or ".param" in l or ".entry" in l or "cvta" in l:
        print(l.strip())
print(type(d), d.__dict__ if hasattr(d, "__dict__") else d)

import os
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/triton_tma.cubin", "wb").write(h.asm["cubin"])
open("gpurun_out/triton_tma.ptx", "w").write(ptx)
