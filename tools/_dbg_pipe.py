import datetime, faulthandler, os, sys, time
faulthandler.dump_traceback_later(45, repeat=True)
from pathlib import Path
from types import SimpleNamespace
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, torch.distributed as dist
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device(f"cuda:{local}"); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=40))
from distributed_pytorch_cookbook_b200.parallel.pipe import PipeStrategy
from distributed_pytorch_cookbook_b200.trainer import build_model
CFG = dict(dim=256, head_dim=32, heads=8, num_layers=4, sequence_length=128, vocab_size=8192, batch_size=8, learning_rate=1e-3)
which = sys.argv[1:] or ["gpipe:never", "1f1b:except_last", "gpipe:except_last"]
for spec in which:
    schedule, ckpt = spec.split(":")
    torch.manual_seed(0)
    ns = SimpleNamespace(**CFG, disable_amp=False, disable_cuda_graph=False, bucket_mb=1.0, comm="auto")
    st = PipeStrategy(num_stages=world, num_microbatches=4, schedule=schedule, dp_size=1, checkpoint=ckpt)
    print(f"[r{rank}] setup {spec}", flush=True)
    st.setup(build_model(ns, CFG["vocab_size"]), ns, dev)
    print(f"[r{rank}] setup done", flush=True)
    g = torch.Generator().manual_seed(3)
    ids_all = torch.randint(0, CFG["vocab_size"], (6, 8, 127), generator=g)
    pos = torch.arange(127).unsqueeze(0).repeat(8, 1).to(dev)
    for s in range(6):
        ids = ids_all[s].to(dev)
        t0 = time.time()
        l = st.train_step({"input_ids": ids, "position_ids": pos}, ids, amp=True)
        torch.cuda.synchronize()
        print(f"[r{rank}] {spec} step {s} loss {float(l):.4f} {1e3*(time.time()-t0):.0f} ms", flush=True)
    dist.barrier()
    st.close()
dist.destroy_process_group()
