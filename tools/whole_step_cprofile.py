"""cProfile (main thread) of three whole-model fine-tune steps: where the host time of the composable encoder nodes goes.
Usage (GPU box): python tools/whole_step_cprofile.py"""
import os, sys, cProfile, pstats, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import m3ae_b200
from m3ae_b200.m3ae_module import M3AETransformerSS
from oracle import m3ae_oracle as O
dev = torch.device("cuda")
torch.manual_seed(0)
model = M3AETransformerSS({"drop_rate": 0.1}, compute_dtype=torch.bfloat16).to(dev).train()
opt = torch.optim.AdamW(list(model.parameters()), lr=1e-5, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01, fused=True)
batches = [O.make_model_batch(70 + i, 64, device=dev) for i in range(2)]
def step(i):
    opt.zero_grad(set_to_none=True)
    loss = model.training_step(batches[i % 2], i)
    loss.backward()
    opt.step()
for i in range(3): step(i)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for i in range(3): step(i)
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(38)
print(s.getvalue()[:9000])
