"""Run under torchrun with >= 2 GPUs: trains each replica for a few steps, averages through the CUDA path
(the peer-memory kernel, or with SPARKNET_AVG=nccl the NCCL all-gather + rank-ordered reduce kernel) and checks the result BIT-EXACTLY against the sequential
host fold (((w0+w1)+w2)+...)/N of the gathered pre-average weights."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from sparknet_b200 import zoo
from sparknet_b200.library import load_library
from sparknet_b200.net import CaffeNet, MinibatchSource
from sparknet_b200.parallel import make_averager

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = load_library(); lib.set_device(local); lib.set_mode_gpu()
name = sys.argv[1] if len(sys.argv) > 1 else "cifar10_quick"
c, h, w, classes, batch, _ = zoo.GEOMETRY[name]
batch = min(batch, 32)
net = CaffeNet.from_solver(zoo.solver_with_net(name, train_bs=batch, test_bs=1, random_seed=1701 + rank, display=0), lib)
rng = np.random.default_rng(1701 + rank)
imgs = [np.ascontiguousarray(rng.random((batch, c, h, w), dtype=np.float32) * 255 - 128) for _ in range(2)]
labs = [rng.integers(0, classes, size=(batch,)).astype(np.float32) for _ in range(2)]
net.setTrainData(MinibatchSource(imgs, labs))
avg = make_averager(lib, net, rank, world)
avg.broadcast_from_rank0()
ok = True
for rnd in range(2):
    net.train(3)
    n = int(lib.sparknet_num_params(net.state))
    mine = np.empty(n, dtype=np.float32)
    lib.sparknet_get_weights_flat(net.state, mine.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n)
    allw = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(world)]
    dist.all_gather(allw, torch.from_numpy(mine).cuda())
    acc = allw[0].cpu().numpy().copy()
    for r in range(1, world):
        acc = (acc + allw[r].cpu().numpy()).astype(np.float32)
    want = (acc / np.float32(world)).astype(np.float32)
    ms = avg.average()
    got = np.empty(n, dtype=np.float32)
    lib.sparknet_get_weights_flat(net.state, got.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n)
    same = got.tobytes() == want.tobytes()
    ok &= same
    if rank == 0:
        print(f"round {rnd}: {n} params, averaging {ms:.3f} ms, bit-exact vs sequential rank-order fold: {same}; "
              f"replicas differed before averaging: {not np.array_equal(allw[0].cpu().numpy(), allw[-1].cpu().numpy())}", flush=True)
if hasattr(avg, "check"):
    avg.check()
if rank == 0:
    print("method:", avg.method, flush=True)
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTIGPU_AVERAGE_OK" if int(flag) == 1 else "MULTIGPU_AVERAGE_MISMATCH", flush=True)
net.close()
dist.destroy_process_group()
sys.exit(0 if int(flag) == 1 else 1)
