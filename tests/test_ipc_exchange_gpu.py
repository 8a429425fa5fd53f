"""IpcPairExchange (copy-engine pull over CUDA IPC + stream memory operations) under
OverlappedBoundary: two ranks, each packs its own windows and must end up with the dequantised
states of its PARTNER's windows, bit-identical to a plain single-process run.  The two ranks share
cuda:0 when the box has one GPU (IPC works between processes on one device), so the test runs in
the 1-GPU tier too.  torch.distributed (gloo) only carries the handle swap."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))
import splitq
from splitq import _native as N
from splitq.transport import IpcPairExchange

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", rank % torch.cuda.device_count())
torch.cuda.set_device(dev)
dist.init_process_group("gloo")
B, H, KVH, S, D, STEPS = 2, 4, 2, 384, 256, int(os.environ.get("SPLITQ_TEST_STEPS", "7"))
FREE_RUNNING = os.environ.get("SPLITQ_TEST_FREE_RUNNING") == "1"
ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)

def inputs(r, step):
    g = torch.Generator().manual_seed(1000 * r + step)
    return (torch.randn(B, H, S, 64, generator=g).to(dev), torch.randn(B, KVH, S, 64, generator=g).to(dev),
            torch.randn(B, S, D, generator=g).to(dev))

def make_plan():
    return splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=N.SCALE_TOKEN, src_device=dev, KVH=KVH)

ob = splitq.OverlappedBoundary(make_plan, depth=2, fused=True)
ipc = IpcPairExchange(ob.plans, rank, world)
for p in ob.plans:
    p.transport = ipc
got = []
if FREE_RUNNING:
    # no host synchronisation between steps: the host runs ahead of both GPUs' streams, slots are
    # reused at depth 2, and each result is copied out by stream-ordered work on the side stream
    ins = [inputs(rank, step) for step in range(STEPS)]
    keep = torch.zeros(STEPS, B, S, D, device=dev)
    torch.cuda.synchronize()
    dist.barrier()
    for step in range(STEPS):
        plan = ob.step(*ins[step])
        with torch.cuda.stream(ob.side):
            keep[step].copy_(plan.h_out)
    ob.drain()
    got = [keep[step] for step in range(STEPS)]
else:
    for step in range(STEPS):
        plan = ob.step(*inputs(rank, step))
        ob.side.synchronize()                      # this step's dequantise is done; keep a copy
        got.append(plan.h_out.clone())
    ob.drain()
ipc.close()
# what a single process computes for the PARTNER's windows
ref = make_plan()
for step in range(STEPS):
    want = ref.run_fused(*inputs(rank ^ 1, step)).hidden
    torch.cuda.synchronize()
    assert torch.equal(got[step].view(torch.int32), want.view(torch.int32)), f"rank {rank} step {step}"
print(f"rank {rank} ok", flush=True)
dist.destroy_process_group()
'''


@pytest.mark.parametrize("free_running", [False, True])
def test_ipc_pair_exchange_two_ranks(tmp_path, free_running):
    """free_running: 50 steps submitted without any host synchronisation (the way bench.py drives the
    exchange), so the two ranks' counters, the depth-2 slot reuse and the copy-engine pulls overlap."""
    script = tmp_path / "worker.py"
    script.write_text(f"ROOT = {ROOT!r}\n" + WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29534" if free_running else "29533", WORLD_SIZE="2",
               PYTORCH_CUDA_ALLOC_CONF="", SPLITQ_TEST_STEPS="50" if free_running else "7",
               SPLITQ_TEST_FREE_RUNNING="1" if free_running else "0")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            pytest.fail("IPC exchange workers timed out")
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"rank {r} ok" in out, out[-3000:]
