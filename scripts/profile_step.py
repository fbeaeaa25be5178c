"""Kernel-time breakdown of one ALBERT-large micro-step (torch.profiler; for analysis only, never for reported numbers)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining
from dedloc_b200.optim.fused import FusedLamb
from dedloc_b200.trainers.pretrain import SyntheticMLMBatches, albert_param_groups

ops = sys.argv[1] if len(sys.argv) > 1 else "native"
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda", 0)
torch.manual_seed(0)
cfg = AlbertConfig.albert_large_v2()
model = AlbertForPreTraining(cfg, impl=ops).to(device=dev, dtype=torch.bfloat16).train()
opt = FusedLamb(albert_param_groups(model, 0.01), lr=1e-3, debias=True)
opt.materialize()
data = SyntheticMLMBatches(cfg.vocab_size, 512, mb, pin=False, pool=2)
batch = data.to_device(data.next_host(), dev)


def step():
    out = model(**batch)
    out.loss.backward()
    model.flush_fused_grads()
    opt.fold_gradients(1.0)


for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    step()
e1.record()
torch.cuda.synchronize()
print(f"micro-step {e0.elapsed_time(e1) / 3:.2f} ms for {mb} samples -> {mb / (e0.elapsed_time(e1) / 3e3):.1f} samples/s")
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=35, max_name_column_width=70))
