"""ncu / timing target: one fast_binless call at chromosome scale (C5: 50 000 bins, band 2000) on a matrix binned on the device.
    python tools/c5_target.py [nbins] [maxdiag] [reads]"""
import sys, time
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
from binless_b200.synth import synth_reads
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
md = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
nr = int(sys.argv[3]) if len(sys.argv) > 3 else 30_000_000
reads = synth_reads(n, 1, 5000, 5, reads_per_dataset=nr)
lib = cuda_lib()
import os
world = int(os.environ.get("WORLD_SIZE", 1))
if world > 1:     # under torchrun: the ranks share the call (trail blocks + shared reduction passes)
    import torch, torch.distributed as dist
    from binless_b200.dist import init_comm
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    init_comm(lib)
h = lib.create_binned(reads, 5000, 1, fetch_mat=False)
for rep in range(int(os.environ.get("C5_REPS", 1))):
  if world > 1:
    dist.barrier()
  t0 = time.perf_counter()
  o = h.run_signal(0.5, 2.5, 0.0, columns="none", maxdiag=md, compute_patchnos=False)
  d = o["diagnostics"]
  if int(os.environ.get("RANK", 0)) == 0:
    print("C5 call", time.perf_counter() - t0, "s; outer steps", d["nsteps"], "solver ms", d["ms_solver"], "launches", h.kernel_launches())
