"""torchrun debug: band-sharded C2 spectrum, step-to-step and vs single GPU."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from transit_b200 import api, workloads, sharding
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
prob = workloads.make_problem("C2")
tr = sharding.BandShardedTransit(lr, rank, world)
tr.ctx.set_stream(sharding.private_stream().cuda_stream)
tr.transit_init(prob)
lo, hi = tr.bands[rank]
first_e = first_s = None
for it in range(8):
    full = tr.do_transit_gathered().cpu().numpy()
    e = tr.ctx.get_extinction()[:, lo:hi]
    loc = tr.device_spectrum_tensor().cpu().numpy()[lo:hi]
    if first_e is None:
        first_e, first_s, first_f = e.copy(), loc.copy(), full.copy()
    de = np.argwhere(e != first_e); ds = np.flatnonzero(loc != first_s); df = np.flatnonzero(full != first_f)
    print(f"rank {rank} it {it}: ext diffs {len(de)} {de[:3].tolist()} local-spectrum diffs {len(ds)} gathered diffs {len(df)} {df[:4].tolist()}", flush=True)
if rank == 0:
    one = api.Transit(lr); one.transit_init(prob); ref = one.do_transit(None)
    d = np.flatnonzero(ref != first_f)
    print("vs single GPU: diffs", len(d), d[:8].tolist(), "bands", tr.bands, flush=True)
dist.barrier(); dist.destroy_process_group()
