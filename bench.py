#!/usr/bin/env python
"""Boundary-path benchmark (BASELINE.json metric: boundary tokens/sec over
score + select + quantise + pack + transfer + dequantise, with HBM roofline of the dominant kernel).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl splitq|reference]

One "step" = one pass of the whole path over a batch of B synthetic evaluation windows.
N > 1 is launched by torchrun (one rank per GPU): every rank processes its own windows and
exchanges its payload with its pair partner (rank^1) over NCCL/NVLink; weak scaling.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200")]

import torch  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[2]: largest single-GPU configuration
    "qwen2-0.5b_split12_colmean_s4096_ladder8-4-2": dict(
        H=14, KVH=2, S=4096, D=896, method="regular_importance", fracs=[(0.25, 2), (0.25, 4), (0.25, 8)],
        scale_mode=1, B=4, B_fused=8),
    # configs[1] shapes (last-row importance, int4 on the lower half, per-token scale)
    "pythia-70m_split3_lastrow_s2048_int4": dict(
        H=8, KVH=8, S=2048, D=512, method="last_row", fracs=[(0.5, 4)], scale_mode=1, B=16, B_fused=32),
    # configs[0] shapes, reference-exact quantiser (one scale for the selected set)
    "pythia-70m_split3_colmean_s512_int4group": dict(
        H=8, KVH=8, S=512, D=512, method="regular_importance", fracs=[(0.5, 4)], scale_mode=0, B=64, B_fused=128),
}
DEFAULT_WORKLOAD = "qwen2-0.5b_split12_colmean_s4096_ladder8-4-2"
METRIC = "boundary tokens/sec (score+quant+pack+xfer+dequant)"
UNIT = "tokens/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def captured_traffic(summary_csv, kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, summed over the launches of one call,
    from a committed `ncu --set full` summary (profiles/); None when the file is not there."""
    import csv
    path = os.path.join(ROOT, "profiles", summary_csv)
    if not os.path.exists(path):
        return None
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    total = 0.0
    for r in rows[2:]:
        if kernel_substr not in r[hdr.index("Kernel Name")]:
            continue
        for col in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(col)
            total += float(r[i]) * scale.get(units[i], 1.0)
    return total or None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index: int, period: float = 0.02) -> None:
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self) -> None:
        if not self.ok:
            return
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def finish(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no samples"}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def synth_inputs(cfg, B, device, seed):
    """Seeded synthetic windows: causal attention probabilities with trained-like peakiness and
    N(0,1) hidden states with 1% outlier channels."""
    H, S, D = cfg["H"], cfg["S"], cfg["D"]
    gen = torch.Generator(device=device).manual_seed(seed)
    mask = torch.full((S, S), float("-inf"), device=device).triu(1)
    A = torch.empty(B, H, S, S, dtype=torch.float32, device=device)
    for b in range(B):
        logits = torch.randn(H, S, S, device=device, generator=gen) * 2.0 + mask
        A[b] = torch.softmax(logits, dim=-1)
        del logits
    h = torch.randn(B, S, D, device=device, generator=gen)
    ch = torch.randperm(D, device=device, generator=gen)[: max(1, D // 100)]
    h[:, :, ch] *= 20.0
    return A, h


def synth_qk(cfg, B, device, seed):
    """Post-RoPE-like Q (B,H,S,64) / K (B,KVH,S,64): N(0,1) entries give logits with unit variance
    after the 1/8 scaling, i.e. trained-like (peaky) attention rows."""
    gen = torch.Generator(device=device).manual_seed(seed + 77)
    q = torch.randn(B, cfg["H"], cfg["S"], 64, device=device, generator=gen)
    k = torch.randn(B, cfg["KVH"], cfg["S"], 64, device=device, generator=gen)
    return q, k


def side_kernel_rooflines(splitq, N, cfg, B, S, D, ladder, h, device):
    """HBM rooflines of the kernels that run beside the scoring (SURVEY 8d algorithmic bytes): each is
    timed ALONE here, after the step loop (in the step they share the GPU with the scoring kernel, so
    their in-step times say nothing about the kernels themselves): CUDA events around 16 back-to-back
    launches, captured into one CUDA graph, that rotate through 4 separate buffer sets (640 MB, so no
    launch finds its data in the 126 MB L2), median of 7 such groups, time per launch."""
    import statistics as st
    peak, peak_src = load_peaks()
    mode = cfg["scale_mode"]
    nbytes = splitq.payload_bytes(S, D, ladder, mode)
    gen = torch.Generator(device=device).manual_seed(99)
    SETS, REP = 4, 16
    hs = [h] + [torch.randn(B, S, D, device=device, generator=gen) for _ in range(SETS - 1)]
    imps = [torch.rand(B, S, device=device, generator=gen) for _ in range(SETS)]
    maps = [splitq.select_bits(imp, ladder, D) for imp in imps]
    payloads = [torch.empty(B, nbytes, dtype=torch.uint8, device=device) for _ in range(SETS)]
    payloads2 = [torch.empty(B, nbytes, dtype=torch.uint8, device=device) for _ in range(SETS)]
    outs = [torch.empty(B, S, D, device=device) for _ in range(SETS)]
    for i in range(SETS):
        splitq.quant_pack(hs[i], maps[i][0], maps[i][1], mode, nbytes, payloads[i])
    group_extra = 0 if mode == N.SCALE_TOKEN else 4 * sum(k for k, _ in ladder) * D * B  # group mode re-reads the selected rows
    cases = [
        ("select_bits", lambda i: splitq.select_bits(imps[i], ladder, D, maps[i][0], maps[i][1]), 4 * B * S + 5 * B * S,
         "latency (S floats per window, one CTA each)"),
        ("quant_pack", lambda i: splitq.quant_pack(hs[i], maps[i][0], maps[i][1], mode, nbytes, payloads[i]),
         4 * B * S * D + 5 * B * S + B * nbytes + group_extra, "hbm"),
        ("transfer_copy", lambda i: splitq.transfer(payloads2[i], payloads[i]), 2 * B * nbytes,
         "hbm (device copy standing in for the hop at N = 1)"),
        ("dequant", lambda i: splitq.dequant(payloads[i], S, D, outs[i]), B * nbytes + 4 * B * S * D, "hbm"),
    ]
    res = {"peak": peak, "peak_source": peak_src, "unit": "GB/s",
           "method": f"each kernel alone: CUDA events around {REP} back-to-back launches (replayed as one CUDA graph) rotating "
                     f"through {SETS} buffer sets (working set 640 MB >> L2), median of 7 groups, per launch"}
    graphed = True
    for name, fn, alg, bound in cases:
        # the REP launches are captured into one CUDA graph and replayed, so that the host's submit time (which is
        # of the order of these kernels' run time) is not part of what the events bracket
        run = None
        try:
            if name == "transfer_copy":   # a copy-engine transfer, not a kernel: as a graph node it is slower than in the step
                raise RuntimeError("not captured")
            for j in range(SETS):
                fn(j)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for j in range(REP):
                    fn(j % SETS)
            run = graph.replay
        except Exception:  # capture refused (e.g. an older driver): time plain launches instead
            torch.cuda.synchronize()
            graphed = graphed and name == "transfer_copy"

            def run(fn=fn):
                for j in range(REP):
                    fn(j % SETS)
        ts = []
        for rep in range(9):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            run()
            b.record()
            torch.cuda.synchronize()
            if rep >= 2:
                ts.append(a.elapsed_time(b) / REP)
        ms = st.median(ts)
        res[name] = {"bound": bound, "kernel_us": 1e3 * ms, "algorithmic_bytes": alg, "achieved": alg / ms / 1e6,
                     "frac": alg / ms / 1e6 / peak, "frac_of_8tbs_spec": alg / ms / 1e6 / 8000.0}
    res["launches"] = "one replayed CUDA graph per group (the copy: plain cudaMemcpyAsync calls)" if graphed else "plain launches"
    return res


def cpu_pass(O, cfg, ladder, fused, inputs):
    """One pass of the reference's CPU path over ONE window (oracle restatement, vectorised ATen)."""
    if fused:
        q, k, h = inputs
        return O.reference_boundary_cpu_from_qk(q, k, h.clone(), 64 ** -0.5, ladder, cfg["scale_mode"], cfg["method"])
    A, h = inputs
    return O.reference_boundary_cpu(A, h.clone(), ladder, cfg["scale_mode"], cfg["method"])


def cpu_reference_leg(cfg, ladder, fused, inputs, min_seconds, threads):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import splitq_oracle as O

    torch.set_num_threads(threads)
    cpu_pass(O, cfg, ladder, fused, inputs)  # warm-up
    times = []
    t_end = time.perf_counter() + min_seconds
    while len(times) < 3 or time.perf_counter() < t_end:
        t0 = time.perf_counter()
        cpu_pass(O, cfg, ladder, fused, inputs)
        times.append(time.perf_counter() - t0)
        if len(times) >= 200:
            break
    return times


def cpu_inputs(cfg, fused, seed=1234):
    S = cfg["S"]
    gen = torch.Generator().manual_seed(seed)
    h = torch.randn(1, S, cfg["D"], generator=gen)
    if fused:
        q = torch.randn(1, cfg["H"], S, 64, generator=gen)
        k = torch.randn(1, cfg["KVH"], S, 64, generator=gen)
        return q, k, h
    mask = torch.full((S, S), float("-inf")).triu(1)
    A = torch.softmax(torch.randn(1, cfg["H"], S, S, generator=gen) * 2.0 + mask, dim=-1)
    return A, h


EVAL_SWEEPS = {
    "qwen": dict(family="qwen2", ratios=[0, 0.25, 0.5, 0.75, 1], layers_of_interest=[3, 11, 18, 22, 23],
                 methods=["regular_importance", "last_row", "aggregate_till"], max_length=512, stride=32),
    "pythia": dict(family="gpt_neox", ratios=[0, 0.25, 0.5, 0.75, 1], layers_of_interest=[1, 2, 4, 5],
                   methods=["regular_importance", "last_row", "aggregate_till"], max_length=2048, stride=32),
}


def run_reference_eval(args):
    """--impl reference --eval-family F: the reference's whole per-window sweep (eager forward for the
    attention maps, get_importance_order, one full layer-by-layer forward per (method, layer, ratio))
    as restated in oracle/, on the host CPU; the CPU leg of benchmarks/eval_throughput.py."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import splitq_oracle as O
    from splitq import compat

    sw = EVAL_SWEEPS[args.eval_family]
    cfg = compat.QWEN2_05B if args.eval_family == "qwen" else compat.PYTHIA_70M
    S, stride = sw["max_length"], sw["stride"]
    model = compat.build_causal_lm(sw["family"], cfg, seed=0)
    eager = compat.clone_with_attention(model, "eager")
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    ids = torch.randint(0, cfg["vocab_size"], (1, S + args.steps * stride), generator=torch.Generator().manual_seed(1234))
    t0 = time.perf_counter()
    for w in range(args.steps):
        win = ids[:, w * stride:w * stride + S]
        O.reference_window_sweep_cpu(model, eager, win, win.clone(), sw["methods"], sw["layers_of_interest"], sw["ratios"])
    dt = time.perf_counter() - t0
    combos = len(sw["methods"]) * len(sw["layers_of_interest"]) * len(sw["ratios"])
    print(json.dumps({"impl": "reference", "metric": "eval window tokens/sec", "eval_family": args.eval_family,
                      "cores": threads, "windows": args.steps, "S": S, "combos_per_window": combos, "seconds": dt,
                      "window_tokens_per_s": args.steps * S / dt,
                      "split_forward_tokens_per_s": args.steps * S * combos / dt}))


def run_eval(args):
    """--eval-family F (product arm): BASELINE.json's second metric, split-inference eval tokens/sec.  The
    reference's sliding-window sweep (methods x split layers x ratios per window) through splitq.evalloop:
    on one GPU both halves run there; on N >= 2 GPUs in the GPU-PAIR layout of SURVEY 8e -- rank 2p runs
    scoring + the upstream half + select/quantise/pack on its GPU, rank 2p+1 dequantise + the downstream
    half + loss on its own, payloads cross by NCCL send/recv, pair p of N/2 owns windows k = p (mod N/2),
    NLL sums are all-reduced.  --steps = windows per pair (after --warmup windows per pair, untimed)."""
    import splitq  # noqa: F401
    from splitq import compat, evalloop
    from splitq.models import Pythia70Model, QwenPointFiveBModel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("TORCH_NCCL_SHOW_EAGER_INIT_P2P_SERIALIZATION_WARNING", "false")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    sw = dict(EVAL_SWEEPS[args.eval_family])
    family = sw.pop("family")
    pairs = world >= 2
    shards = world // 2 if pairs else 1
    S, stride = sw["max_length"], sw["stride"]
    if family == "qwen2":
        cfg, wrapper = compat.QWEN2_05B, None
        wrapper = QwenPointFiveBModel.from_config(dev, cfg=cfg, seed=0)
    else:
        cfg = compat.PYTHIA_70M
        wrapper = Pythia70Model.from_config(dev, sw["ratios"], cfg=cfg, seed=0)
    fn = evalloop.make_importance_fn(wrapper, sw["methods"], "fused", layers=sw["layers_of_interest"])

    def sweep(windows):
        n = S + (windows - 1) * stride
        ids = torch.randint(0, cfg["vocab_size"], (1, n), generator=torch.Generator().manual_seed(1234))
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = evalloop.evaluate(sw, wrapper, ids, fn, S, reduce_device=dev if dist is not None else None,
                                log=lambda *a: None, batch_combos=True, pairs=pairs)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = t.item()
        return res, dt

    sweep(max(1, args.warmup) * shards)
    windows = args.steps * shards
    res, dt = sweep(windows)
    combos = len(sw["methods"]) * len(sw["layers_of_interest"]) * len(sw["ratios"])
    if rank == 0:
        print(json.dumps({
            "metric": "eval window tokens/sec (split-inference sweep)", "value": windows * S / dt, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"eval-{args.eval_family}", "model": family, "S": S, "stride": stride,
                       "combos_per_window": combos, "windows": windows, "windows_scored": res.windows,
                       "layout": (f"{shards} GPU pair(s): upstream half on rank 2p, downstream half on rank 2p+1, payload over "
                                  "NCCL send/recv" if pairs else "both halves on one GPU"),
                       "batching": "all (method, ratio > 0) variants of a split layer as one downstream batch"},
            "split_forward_tokens_per_s": windows * S * combos / dt, "scored_tokens": res.n_tokens}))
    if dist is not None:
        dist.destroy_process_group()


def run_reference(args, cfg, name):
    """--impl reference: the reference's CPU implementation of the same path, all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    S = cfg["S"]
    fused = args.importance == "fused"
    ladder = [(int(f * S), b) for f, b in cfg["fracs"]]
    threads = os.cpu_count() or 1
    inputs = cpu_inputs(cfg, fused)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import splitq_oracle as O

    torch.set_num_threads(threads)
    for _ in range(args.warmup):
        cpu_pass(O, cfg, ladder, fused, inputs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(O, cfg, ladder, fused, inputs)
    dt = time.perf_counter() - t0
    value = args.steps * S / dt
    start = "Q,K (eager attention probabilities computed on the CPU, as HF eager does)" if fused else "materialised A"
    sample = (f"{args.steps} steps x 1 window of {S} tokens (H={cfg['H']}, D={cfg['D']}) from {start}, "
              f"torch CPU fp32")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": name, "windows_per_step": 1, "importance_impl": args.importance},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="splitq", choices=["splitq", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--importance", default="fused", choices=["fused", "materialised"],
                    help="fused: scores from the split layer's Q,K (K2, no attention matrix); "
                         "materialised: scores from attention probabilities already in HBM (K1)")
    ap.add_argument("--fused-mode", default="f16x3", choices=["f16x3", "f16x1"])
    ap.add_argument("--transfer", default="auto", choices=["auto", "nccl", "ipc"],
                    help="payload hop between paired ranks in the device-resident loop: NCCL send/recv, or a "
                         "copy-engine pull over CUDA IPC ordered by stream memory operations (auto = ipc, "
                         "falling back to nccl if the mapping cannot be set up)")
    ap.add_argument("--main-priority", type=int, default=1,
                    help="> 0 (default): score on the caller's stream.  <= 0: score on a separate stream of that CUDA "
                         "priority (negative = above the side stream's 0); measured in round 2: the per-step join of "
                         "that stream with the caller's costs 30 us of idle scoring stream per step on one GPU and "
                         "100 us across GPUs, more than the priority gains (profiles/r2_d_stream_variants.txt)")
    ap.add_argument("--pack-on-main", action="store_true",
                    help="keep select + quantise/pack on the scoring stream (default: on the side stream with hop + dequantise)")
    ap.add_argument("--batch-windows", type=int, default=0, help="windows per step (0 = workload default)")
    ap.add_argument("--depth", type=int, default=0,
                    help="buffer sets of the software pipeline (0 = 2 on one GPU, 3 across GPUs: the third set lets the "
                         "NVLink pull of step i run under the scoring of step i+1 instead of between two scorings)")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eval-family", default=None, choices=["qwen", "pythia"],
                    help="time the sliding-window eval sweep instead of the boundary path: --steps windows (per GPU pair); "
                         "with --impl reference the CPU port of the reference loop")
    args = ap.parse_args()
    cfg = WORKLOADS[args.workload]
    if args.impl == "reference" and args.eval_family:
        run_reference_eval(args)
        return
    if args.impl == "reference":
        run_reference(args, cfg, args.workload)
        return
    if args.eval_family:
        run_eval(args)
        return
    args.warmup = max(args.warmup, 3)

    import splitq
    from splitq import _native as N
    from splitq.transport import PairExchange

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl splitq needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torchrun)"

    fused = args.importance == "fused"
    B = args.batch_windows or (cfg["B_fused"] if fused else cfg["B"])
    H, KVH, S, D = cfg["H"], cfg["KVH"], cfg["S"], cfg["D"]
    ladder = splitq.ladder_from_fractions(cfg["fracs"], S)
    plan = splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=cfg["scale_mode"], method=cfg["method"],
                               src_device=device, KVH=KVH,
                               fused_mode=N.FUSED_F16X3 if args.fused_mode == "f16x3" else N.FUSED_F16X1)
    plan.transport = PairExchange(rank, world) if world > 1 else None
    if fused:
        q, k = synth_qk(cfg, B, device, seed=1234 + rank)
        gen = torch.Generator(device=device).manual_seed(1234 + rank)
        h = torch.randn(B, S, D, device=device, generator=gen)
        h[:, :, torch.randperm(D, device=device, generator=gen)[: max(1, D // 100)]] *= 20.0
        A = None
    else:
        A, h = synth_inputs(cfg, B, device, seed=1234 + rank)

    def make_plan():
        pl = splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=cfg["scale_mode"], method=cfg["method"],
                                 src_device=device, KVH=KVH, fused_mode=plan.fused_mode)
        pl.transport = plan.transport
        return pl

    # software pipeline over steps: hop + dequantise of step i on a side stream under the scoring of i+1
    depth = args.depth or (2 if world == 1 else 3)
    ob = splitq.OverlappedBoundary(make_plan, depth=depth, fused=fused, pack_on_side=not args.pack_on_main,
                                   main_priority=None if args.main_priority > 0 else args.main_priority)
    dev_inputs = (q, k, h) if fused else (A, h)
    ipc, transfer_name = None, "nccl pair exchange" if world > 1 else "device-to-device copy"
    if world > 1 and args.transfer in ("auto", "ipc"):
        from splitq.transport import IpcPairExchange
        try:  # the constructor fails on all ranks or on none
            ipc = IpcPairExchange(ob.plans, rank, world)
            for pl in ob.plans:
                pl.transport = ipc
            transfer_name = "copy-engine pull over CUDA IPC (NVLink P2P) ordered by stream memory operations; no SM"
        except N.SplitqError as e:  # e.g. a virtual-memory allocator, no peer access
            if args.transfer == "ipc":
                raise
            print(f"[bench] rank {rank}: {e}; using NCCL", file=sys.stderr)
            ipc = None

    def step(timed_events=None, side_event=None):
        ob.step(*dev_inputs, score_events=timed_events, side_event=side_event)

    sampler = ClockSampler(local)  # NVML set-up takes milliseconds and differs per rank: keep it out of the start line
    for _ in range(args.warmup):
        step()
    ob.drain()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    sampler.start()
    if dist is not None:
        # device-side start line: the ranks leave the host barrier up to milliseconds apart, and a rank that
        # starts early only waits for its partner's first payload (round 1: one such wait halved the 8-GPU
        # number of a 16 ms timed region).  This all-reduce is queued on the scoring stream right before the
        # start event, so every rank's GPU begins step 0 when the slowest rank's host has got here.
        start_line = torch.zeros(1, device=device)
        dist.all_reduce(start_line)
    k_events = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    side_events = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    host_t = [0.0] * (args.steps + 1)
    if ipc is not None:
        ipc.trace = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(args.steps):
        host_t[i] = time.perf_counter()
        step(k_events[i], side_events[i])
    host_t[args.steps] = time.perf_counter()
    ob.join()  # the clock stops when the last step's dequantise has finished
    t1.record()
    torch.cuda.synchronize()
    host_done = time.perf_counter()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    elapsed_ms = t0.elapsed_time(t1)
    score_ms = sum(a.elapsed_time(b) for a, b in k_events) / args.steps
    # where each rank's time went, step by step (all from CUDA events on the streams the work ran on,
    # plus the host's submit clock): a step whose completion interval is far above the median names
    # the stall; `main_gap` = scoring stream idle between two scorings (waits for a free slot),
    # `partner_wait` = side stream held by splitq_stream_wait_geq_u32 until the partner's payload is
    # complete, `host_submit` = wall time the host needed to enqueue one step (>= the step time
    # means the GPU is starved by the launch path, not by kernels)
    done_at = [t0.elapsed_time(e) for e in side_events]
    step_ms = [done_at[0]] + [done_at[i] - done_at[i - 1] for i in range(1, args.steps)]
    score_each = [a.elapsed_time(b) for a, b in k_events]
    main_gap = [k_events[i - 1][1].elapsed_time(k_events[i][0]) for i in range(1, args.steps)] or [0.0]
    host_submit = [1e3 * (host_t[i + 1] - host_t[i]) for i in range(args.steps)]
    partner_wait = [a.elapsed_time(b) for a, b in (ipc.trace if ipc is not None else [])][-args.steps:] or [0.0]
    med = statistics.median
    per_rank = {
        "rank": rank, "total_ms": elapsed_ms,
        "step_ms": {"min": min(step_ms), "median": med(step_ms), "max": max(step_ms), "slowest_step": step_ms.index(max(step_ms))},
        "score_ms": {"median": med(score_each), "max": max(score_each), "slowest_step": score_each.index(max(score_each))},
        "main_gap_ms": {"median": med(main_gap), "max": max(main_gap), "sum": sum(main_gap)},
        "partner_wait_ms": {"median": med(partner_wait), "max": max(partner_wait), "sum": sum(partner_wait)},
        "host_submit_ms": {"median": med(host_submit), "max": max(host_submit), "sum": sum(host_submit),
                           "host_ran_ahead_ms": 1e3 * (host_done - host_t[args.steps])},
    }
    if dist is not None:
        gathered = [None] * world
        dist.all_gather_object(gathered, per_rank)
        per_rank_all = gathered
    else:
        per_rank_all = [per_rank]
    if ipc is not None:
        ipc.close()

    # ---- end to end through the public host-buffer API (splitq.HostPipeline): every step uploads its
    # inputs from pinned host memory, runs the path, and downloads the dequantised hidden states;
    # consecutive steps overlap on three streams, the clock stops when the last result is on the host
    pin = lambda t: torch.empty(t.shape, dtype=torch.float32, pin_memory=True).copy_(t)  # noqa: E731
    h_host, out_host = pin(h), torch.empty(h.shape, dtype=torch.float32, pin_memory=True)
    inputs_host = [pin(q), pin(k), h_host] if fused else [pin(A), h_host]

    pipe = splitq.HostPipeline(make_plan, depth=2, fused=fused)
    torch.cuda.synchronize()
    for _ in range(2):
        pipe.submit(inputs_host, out_host)
    pipe.drain()
    if dist is not None:
        dist.barrier()
    e0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        pipe.submit(inputs_host, out_host)
    pipe.drain()
    e2e_s = time.perf_counter() - e0
    clocks = sampler.finish()

    if dist is not None:
        t = torch.tensor([elapsed_ms, e2e_s, score_ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms, e2e_s, score_ms = t.tolist()

    aux = side_kernel_rooflines(splitq, N, cfg, B, S, D, ladder, h, device) if rank == 0 else None
    if rank == 0:
        peaks = {}
        path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(path):
            peaks = json.load(open(path))
        tokens_per_step = world * B * S
        value = tokens_per_step * args.steps / (elapsed_ms * 1e-3)
        if fused and cfg["method"] == "last_row":
            # last-row importance from Q,K is one GEMV per head: HBM-bound, no tensor-core work
            peak, peak_src = load_peaks()
            alg_bytes = 4 * B * (KVH * S * 64 + H * 64 + S)
            achieved = alg_bytes / (score_ms * 1e-3) / 1e9
            roofline = {"bound": "hbm", "kernel": "last_row_logits_kernel + last_row_softmax_kernel", "achieved": achieved,
                        "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": score_ms,
                        "note": "tiny, latency-bound reduction; the step is dominated by quantise/pack + dequantise"}
            l2_note = f"inputs per step ({4 * B * S * (64 * (H + KVH) + D) / 1e6:.0f} MB) exceed the 126 MB L2; no flush"
        elif fused:
            nb = (S + 127) // 128
            tn = N.load().splitq_importance_fused_tile_cols()    # a tile = 128 fixed rows x tn streamed rows
            nbn = (S + tn - 1) // tn
            # causal tiles of the two passes (the formulas of importance_fused.cu): pass 1 = query blocks x key
            # tiles up to the block's last query, pass 2 = key blocks x query tiles from the block's first key on
            tiles = B * H * (sum((min((b + 1) * 128, S) + tn - 1) // tn for b in range(nb)) +
                             sum(nbn - (b * 128) // tn for b in range(nb)))
            tile_flops = 2.0 * 128 * tn * 64
            # the kernel is timed inside a long back-to-back step loop: the sustained cuBLAS figure is its peak
            bf16_key = "bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else "bf16_tflops"
            bf16 = peaks.get(bf16_key, 1590.0)
            x3 = args.fused_mode == "f16x3"
            issued = tiles * tile_flops * (3 if x3 else 1)        # h0*h0 (+ h1*h0 + h0*h1), all kind::f16
            peak_mix = bf16                                       # kind::f16 runs at the measured cuBLAS bf16 rate
            dense_equiv = 4.0 * B * H * S * S * 64
            achieved = issued / (score_ms * 1e-3) / 1e12
            mufu_ms = tiles * 128 * tn / (16.0 * 148 * 1.965e9) * 1e3   # 16 ex2/clk/SM at the max SM clock
            burst = peaks.get("bf16_tflops", 1590.0)
            roofline = {
                "bound": "tensor", "kernel": "fused_importance_kernel (2 passes) + operand split + lse merge + finalize",
                "achieved": achieved, "peak": peak_mix, "unit": "TFLOP/s", "frac": achieved / peak_mix,
                "frac_of_sustained_peak": achieved / peak_mix, "frac_of_burst_peak": achieved / burst,
                "which_peak_applies": "sustained: the kernel runs back to back inside the step loop and the SM clock sits "
                                      "near 1.65 GHz under sw_power_cap while it does (ncu: cycles / duration)",
                "traffic": (captured_traffic("r2_i_ncu_full_k2_summary.csv", "fused_importance_kernel")
                            if args.workload == DEFAULT_WORKLOAD and B == cfg["B_fused"] and x3 else None),
                "traffic_note": "DRAM bytes of the two tensor-core passes (ncu); operands are re-read from L2, not HBM",
                "peak_source": (f"MEASURED_PEAKS.json {bf16_key}" if peaks else "fallback 1590 (B200_PROFILING.md)") +
                               " (kind::f16 runs at the cuBLAS bf16 rate; burst figure: " +
                               f"{peaks.get('bf16_tflops', 1590.0):.0f})",
                "flops_counted": f"issued tcgen05 MMA flops: causal 128 x {tn} tiles of both passes x 3 fp16 products (split operands)",
                "useful_128x128_tile_tflops": B * H * nb * (nb + 1) * 2.0 * 128 * 128 * 64 * (3 if x3 else 1) / (score_ms * 1e-3) / 1e12,
                "issued_flops_per_launch": issued, "dense_equivalent_tflops": dense_equiv / (score_ms * 1e-3) / 1e12,
                "kernel_ms": score_ms, "co_bound_mufu_frac": mufu_ms / score_ms,
                "attention_matrix_bytes_avoided": 4 * B * H * S * S}
            l2_note = f"inputs per step ({4 * B * S * (64 * (H + KVH) + D) / 1e6:.0f} MB) exceed the 126 MB L2; no flush"
        else:
            peak, peak_src = load_peaks()
            if cfg["method"] == "last_row":
                alg_bytes, kname = 4 * B * H * S + 4 * B * S, "importance_finalize_kernel (last row)"
            else:
                alg_bytes, kname = 4 * B * H * S * S + 4 * B * S, "colsum_partial_kernel + importance_finalize_kernel"
            achieved = alg_bytes / (score_ms * 1e-3) / 1e9
            traffic = None
            if args.workload == DEFAULT_WORKLOAD and B == cfg["B"]:  # the committed capture is of exactly this launch
                traffic = captured_traffic("r1_a_ncu_full_summary.csv", "colsum_partial_kernel")
            roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "frac_of_8tbs_spec": achieved / 8000.0, "traffic": traffic,
                        "peak_source": peak_src,
                        "traffic_source": "profiles/r1_a_ncu_full_summary.csv (ncu --set full, dram read+write)" if traffic else None,
                        "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": score_ms}
            l2_note = f"inputs per step ({alg_bytes / 1e6:.0f} MB) exceed the 126 MB L2; no flush"
        h2d, d2h = plan.host_bytes_per_run(fused)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "windows_per_step_per_gpu": B, "H": H, "KVH": KVH, "S": S, "D": D,
                       "ladder": ladder, "scale_mode": "token" if cfg["scale_mode"] else "group",
                       "importance_impl": args.importance + (f" ({args.fused_mode})" if fused else ""),
                       "payload_bytes_per_window": plan.nbytes,
                       "transfer": transfer_name, "l2": l2_note,
                       "pipelining": ("hop + dequantise" if args.pack_on_main else "select + quantise/pack + hop + dequantise") +
                                     " of step i on a side stream, overlapping the scoring of step i+1"},
            "roofline": roofline,
            "roofline_aux": aux,
            "e2e": {"value": world * B * S * args.e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": args.e2e_steps,
                    "api": "splitq.HostPipeline.submit/drain (pinned host buffers, 3 streams, depth 2)"},
            "per_rank_ms": per_rank_all,
            "median_step": {"ms": max(r["step_ms"]["median"] for r in per_rank_all),
                            "value": tokens_per_step / (max(r["step_ms"]["median"] for r in per_rank_all) * 1e-3),
                            "note": "tokens/s at the slowest rank's MEDIAN step-completion interval; `value` above "
                                    "stays total tokens / total time"},
            "gpu_launches": (plan.launches_per_run(fused) - (3 if fused and cfg["method"] == "last_row" else 0)) * args.steps,
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            inputs = (q[0:1].cpu(), k[0:1].cpu(), h[0:1].cpu()) if fused else (A[0:1].cpu(), h[0:1].cpu())
            times = cpu_reference_leg(cfg, ladder, fused, inputs, args.cpu_seconds, threads)
            out["cpu_baseline"] = {
                "value": S / statistics.mean(times), "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"{len(times)} passes over 1 window of {S} tokens (same H, D, ladder; starts from "
                          f"{'Q,K' if fused else 'materialised attention'}), torch CPU fp32, "
                          f"mean {1e3 * statistics.mean(times):.1f} ms/window"}
            one = cpu_reference_leg(cfg, ladder, fused, inputs, min(2.0, args.cpu_seconds), 1)
            out["cpu_baseline"]["value_1_thread"] = S / statistics.mean(one)
            torch.set_num_threads(threads)
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
