import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import torch
import finito_b200 as fb
from fin_datasets import make_dataset
ds = make_dataset("bench", genome_len=5_000_000, k=31, n_reads=10_000_000, read_len=150, seed=20260102, families=50)
out = os.path.join(ds["dir"], "probe_build")
def go(tag):
    for i in range(3):
        t0 = time.perf_counter(); fb.build_fmin(ds["sbwt"], ds["unitigs"], out); dt = time.perf_counter() - t0
    t = fb.build_timing()
    print(tag, f"wall {dt*1e3:.1f} ms", {k: round(v, 2) for k, v in t.items() if k != "unitig_bases"}, flush=True)
go("fresh            ")
d = torch.empty(12_000_000_000, dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
go("12 GB device held ")
del d; torch.cuda.synchronize()
go("12 GB cached      ")
torch.cuda.empty_cache()
go("cache emptied     ")
h = torch.empty(9_600_000_000, dtype=torch.uint8).pin_memory()
go("9.6 GB pinned held")
del h
go("pinned freed      ")
ix = fb.FinimizerIndex.load(out)
go("index loaded      ")
