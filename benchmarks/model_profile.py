"""Per-kernel time table (torch.profiler / CUPTI) of a zoo model's training step.
usage: model_profile.py {resnet18|bert_base|bert_large} out.txt"""
import os
import runpy
import sys

import torch

name, out_path = sys.argv[1], sys.argv[2]
here = os.path.dirname(os.path.abspath(__file__))
sys.argv = ['model_step.py', name, '6']
g = runpy.run_path(os.path.join(here, 'model_step.py'))          # builds the executor, runs 6 steps (graph captured)
ex, fd = g['ex'], g['fd']
from torch.profiler import profile, ProfilerActivity               # noqa: E402
steps = 5
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        ex.run(feed_dict=fd)
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in rows)
with open(out_path, 'w') as f:
    f.write('%s training step, %d steps, sum of kernel time per step %.1f us\n' % (name, steps, tot / steps))
    for e in rows[:45]:
        f.write('%9.1f us/step %5.1f%%  x%-5.1f %s\n' % (e.device_time_total / steps, 100.0 * e.device_time_total / tot, e.count / steps, e.key[:120]))
print(open(out_path).read())
