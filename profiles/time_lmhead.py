"""Times the tcgen05 LM-head kernel at the bench shape; AREAL_B200_LMHEAD_DEBUG / _CTA_GROUP select variants."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import lm_head as LM
D = torch.device("cuda:0")
T, Hd, V = int(os.environ.get("PROF_T", 16384)), int(os.environ.get("PROF_H", 1536)), 151936
g = torch.Generator(device=D).manual_seed(3)
h = torch.randn(T, Hd, generator=g, device=D).bfloat16()
w = (torch.randn(V, Hd, generator=g, device=D) * (2.0 / Hd ** 0.5)).bfloat16()
lab = torch.randint(0, V, (T,), generator=g, device=D)
for _ in range(3):
    LM.lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=os.environ.get("PROF_ENT", "1") == "1")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    LM.lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=os.environ.get("PROF_ENT", "1") == "1")
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"debug={os.environ.get('AREAL_B200_LMHEAD_DEBUG','0')} cg={os.environ.get('AREAL_B200_LMHEAD_CTA_GROUP','2')} "
      f"T={T} H={Hd}: {ms:.3f} ms  {2.0*T*Hd*V/ms/1e9:.0f} TFLOP/s")
