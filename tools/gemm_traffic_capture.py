"""Runs the forward of ONE fused text+image BertCrossLayer pair at batch 64 (dropout 0.1, bf16) four times: under
`ncu --set full -k regex:gemm_tcgen05 -s 18 -c 6` the six captured launches are the six GEMM stages of the 4th forward
(QKV, out-proj, cross Q+KV, cross out-proj, FFN-up, FFN-down) with their producers' outputs still in L2, as in the step.
Usage (GPU box): ncu --set full --clock-control none -k regex:gemm_tcgen05 -s 18 -c 6 -f -o gpurun_out/ncu_gemm_traffic python tools/gemm_traffic_capture.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import m3ae_b200  # noqa: F401
from m3ae_b200.bert_cross_layer import BertCrossLayer, fused_layer_pair
from transformers import RobertaConfig

dev = torch.device("cuda")
cfg = RobertaConfig(vocab_size=50265, hidden_size=768, num_hidden_layers=6, num_attention_heads=12, intermediate_size=3072,
                    max_position_embeddings=32, hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1)
torch.manual_seed(0)
lt, li = BertCrossLayer(cfg).to(dev).train(), BertCrossLayer(cfg).to(dev).train()
B = 64
x = torch.randn(B, 32, 768, device=dev).to(torch.bfloat16)
y = torch.randn(B, 325, 768, device=dev).to(torch.bfloat16)
mx = torch.zeros(B, 1, 1, 32, device=dev)
my = torch.zeros(B, 1, 1, 325, device=dev)
with torch.no_grad():
    for _ in range(4):
        (x1,), (y1,) = fused_layer_pair(lt, li, x, y, mx, my)
torch.cuda.synchronize()
print("ok", x1.shape, y1.shape)
