"""The NCCL-path parity check of swift_subtask_dev_b200.multi on its own (bench.py runs the same check before
the timed loop):  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
--master-port P scripts/nccl_parity.py u|c|uc|s   (u: uniform box, c: clustered box in weighted blocks,
s: uniform box in the SPHENIX scheme with its three halo exchanges)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch, torch.distributed as dist
from swift_subtask_dev_b200 import multi
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tr = multi.NcclTransport()
which = sys.argv[1] if len(sys.argv) > 1 else "uc"
rc = 0
try:
    if "u" in which:
        print("rank", rank, multi.nccl_parity_check(tr, world, rank, local), flush=True)
    if "c" in which:
        print("rank", rank, multi.nccl_parity_check(tr, world, rank, local, clustered=True), flush=True)
    if "s" in which:
        print("rank", rank, multi.nccl_parity_check(tr, world, rank, local, sphenix=True), flush=True)
except Exception as e:
    print("rank", rank, "ERR", str(e)[:200], flush=True)
    rc = 1
dist.barrier(); dist.destroy_process_group()
sys.exit(rc)
