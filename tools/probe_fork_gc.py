"""GPU probe: does a forked DataLoader worker that garbage-collects an inherited, orphaned CUDA graph abort?
(plain iter(loader), i.e. WITHOUT _ShardedLoader's collect + freeze)"""
import gc, sys
import torch
from torch.utils.data import DataLoader, Dataset

class DS(Dataset):
    def __len__(self): return 4
    def __getitem__(self, i):
        gc.collect()
        return torch.full((2,), float(i))

class Cycle: pass
dev = torch.device("cuda:0")
x = torch.ones(8, device=dev)
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    y = x * 2
graph.replay(); torch.cuda.synchronize()
c = Cycle(); c.me, c.graph, c.event, c.pinned = c, graph, torch.cuda.Event(), torch.empty(1 << 20).pin_memory()
c.event.record()
gc.collect(); gc.collect()
del c, graph, y
try:
    print("plain iter:", [b.tolist() for b in DataLoader(DS(), batch_size=2, num_workers=2)])
except Exception as e:
    print("plain iter FAILED:", type(e).__name__, str(e)[:200])
