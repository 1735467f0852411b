#!/usr/bin/env python
"""Headline benchmark: target-points/sec, forward + loss + backward (+ gradient all-reduce + Adam).

    python bench.py --gpus N --steps K --warmup W            # product arm (CUDA kernels through the module API)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the reference algorithm (oracle port)

Workload at every N: BASELINE.json configs[1] per GPU -- AttnCNP (transformer attender, r_dim 128) on synthetic
chirp waveforms, 4096 samples, 512 context / 4096 target points, batch 256 tasks per GPU (weak scaling: the
global batch is 256*N, sharded over ranks; the only exchange is the gradient all-reduce).  One "step" =
forward + CNPFLoss + backward + all-reduce + Adam.  Prints ONE JSON line on rank 0.
"""
import argparse
import datetime
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "target-points/sec fwd+bwd"
UNIT = "target-points/s"
WORKLOADS = {
    # name: (model, tasks per GPU, n_context, n_target, n_z)
    "attncnp_c2": ("AttnCNP", 256, 512, 4096, 1),
    "convcnp_c3": ("ConvCNP_unet", 32, 512, 4096, 1),
    "attnlnp_c4": ("AttnLNP", 16, 512, 4096, 32),
    "cnp_c1": ("CNP", 64, 64, 128, 1),
    # BASELINE.json configs[4] (inference sweep, forward only; one resident chunk of the 1M-task sweep per step)
    "attncnp_c5_infer": ("AttnCNP", 64, 512, 16384, 1),
    "convcnp_c5_infer": ("ConvCNP_unet", 64, 512, 16384, 1),
}
WORKLOAD_TEXT = {
    "attncnp_c2": "BASELINE.json configs[1]: AttnCNP (transformer attender, r_dim=128) on synthetic 1D chirp waveforms, "
                  "4096 samples, 512 context / 4096 target points, batch 256 tasks per GPU, training step",
    "convcnp_c3": "BASELINE.json configs[2]: ConvCNP 1D (SetConv + UNet, density 64) on chirps, 512/4096, 32 tasks per GPU",
    "attnlnp_c4": "BASELINE.json configs[3]: AttnLNP, 32 z-samples, ELBO, 512/4096, 16 tasks per GPU",
    "cnp_c1": "BASELINE.json configs[0]: CNP, batch 64, 64 context / 128 target",
    "attncnp_c5_infer": "BASELINE.json configs[4]: batched inference, AttnCNP predicting 16384-sample waveforms from 512 "
                        "context points, chunks of 64 tasks per GPU, no communication",
    "convcnp_c5_infer": "BASELINE.json configs[4]: batched inference, ConvCNP (SetConv + UNet) predicting 16384-sample "
                        "waveforms from 512 context points, chunks of 64 tasks per GPU, no communication",
}
# algorithmic forward FLOPs per target point; backward = 2x
FLOPS_FWD_PER_TARGET = {"attncnp_c2": 616.2e3, "convcnp_c3": 280.7e3, "attnlnp_c4": 8.93e6, "cnp_c1": 316.8e3,
                        "attncnp_c5_infer": 597.0e3, "convcnp_c5_infer": 231.0e3}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="attncnp_c2", choices=list(WORKLOADS))
    ap.add_argument("--mode", default=os.environ.get("NPFK_MODE", "fast"), choices=["fp32", "fast"])
    ap.add_argument("--tasks-per-gpu", type=int, default=0, help="override the per-GPU batch (debugging only)")
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"],
                    help="replay forward+loss+backward as one CUDA graph (auto = on); off = eager launches only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: fixed tasks per GPU; strong: BASELINE.json's global batch split over the ranks")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the secondary block (configs[2] and configs[3] measured after the headline workload)")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="budget of the cpu_baseline leg")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=p["hbm_gbs"], tflops=p["bf16_tflops_sustained"], source="measured (MEASURED_PEAKS.json, sustained)")
    return dict(hbm_gbs=6650.0, tflops=1400.0, source="fallback (B200_PROFILING.md)")


# ------------------------------------------------------------------------------------------ CPU arm / cpu_baseline
def cpu_reference_rate(workload, seconds, warmup=1, min_iters=2, max_iters=50, chunk=None):
    """Times the reference algorithm (oracle port, torch CPU fp32, all host threads) on a bounded chunk of the
    workload: forward + loss + backward.  Returns (target-points/s, description, cores, ms per chunk)."""
    import torch

    from oracle import npf_oracle as O
    from util_models import build_product, chirp_batch

    model, B, Nc, Nt, nz = WORKLOADS[workload]
    infer = workload.endswith("_infer")
    chunk = chunk or {"attncnp_c2": 4, "convcnp_c3": 2, "attnlnp_c4": 1, "cnp_c1": 64}.get(workload, 1)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    sd = {k: v.detach().clone() for k, v in build_product(model, 128, nz=nz).state_dict().items()}
    inp = chirp_batch(chunk, Nc, Nt, 1234)
    args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"])
    kw = {}
    fwd = O.FORWARDS[model.split("_")[0]]
    loss_fn = O.cnpf_loss
    if model.startswith("ConvCNP"):
        kw = dict(unet=model.endswith("unet"), density_induced=64)
    if "LNP" in model:
        args["y_t"] = inp["Y_trgt"]
        kw = dict(n_z=nz)
        loss_fn = O.elbo_loss

    def one():
        if infer:
            with torch.no_grad():
                fwd(sd, **args, training=False, **kw)
        else:
            O.loss_and_grads(fwd, loss_fn, sd, args, inp["Y_trgt"], training=True, **kw)

    for _ in range(warmup):
        one()
    times = []
    t_end = time.time() + seconds
    while (len(times) < min_iters or time.time() < t_end) and len(times) < max_iters:
        t0 = time.time()
        one()
        times.append(time.time() - t0)
    t = statistics.median(times)
    desc = (f"{chunk}-task chunk of the workload ({Nc} context / {Nt} target, n_z={nz}), "
            f"{'forward only (eval)' if infer else 'forward+loss+backward'}, "
            f"median of {len(times)} runs after {warmup} warm-up, torch {torch.__version__} CPU fp32, {threads} threads")
    return chunk * Nt / t, desc, threads, t * 1e3


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_iters = max(args.steps, 1)
    rate, desc, cores, ms = cpu_reference_rate(args.workload, seconds=1e9, warmup=max(args.warmup, 1), min_iters=n_iters,
                                               max_iters=n_iters)
    model, B, Nc, Nt, nz = WORKLOADS[args.workload]
    line = {
        "impl": "reference",
        "metric": "target-points/sec forward (batched inference)" if args.workload.endswith("_infer") else METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "sample": desc, "timing": "host wall clock"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference = unmodified algorithm of MarinerQ/npf_GWwaveform restated in oracle/npf_oracle.py (the Python "
                "reference itself cannot travel to the GPU box); runs on host cores only",
    }
    emit(line)


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, f"/tmp/npf_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as f:
            for ln in f:
                parts = [p.strip() for p in ln.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx.append(float(parts[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, parts[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------ roofline
# kernel family -> (C-ABI entry points timed with CUDA events on the launching stream, roofline that bounds it,
#                   device kernel whose ncu --set full DRAM traffic is recorded in profiles/ncu_traffic.json)
KERNEL_FAMILIES = {
    "attention backward (tcgen05)": (["npfk_mha_bwd"], "tensor", "attn_bwd_tc_kernel"),
    "attention forward (tcgen05)": (["npfk_mha_fwd"], "tensor", "attn_fwd_tc_kernel"),
    "MLP chain forward (K1: layers fused on chip, tcgen05 + TMEM-resident activations)": (
        ["npfk_mlp_chain_fwd"], "hbm", "tc_mlp_chain_fwd_kernel"),
    "decoder tail backward (log-likelihood + head dgrad/wgrad)": (["npfk_nll_head_bwd"], "hbm", "nll_head_bwd_kernel"),
    "linear forward (tcgen05 GEMM + bias/residual/ReLU epilogue)": (["npfk_linear_fwd"], "hbm", "tc_gemm_nt_kernel<0>"),
    "linear backward fused (dgrad + mask + wgrad)": (["npfk_linear_bwd_fused"], "hbm", "tc_linear_bwd_fused_kernel"),
    "linear backward, two layers per launch (K1b: the gradient between them stays on chip)": (
        ["npfk_linear_bwd_fused2"], "hbm", "tc_chain_bwd2_kernel"),
    "linear dgrad": (["npfk_linear_bwd_dx"], "hbm", "tc_gemm_nt_kernel<1>"),
    "linear wgrad": (["npfk_linear_bwd_dw"], "hbm", "tc_gemm_tn_kernel"),
    "residual + LayerNorm": (["npfk_add_layernorm_fwd", "npfk_add_layernorm_bwd"], "hbm", "add_layernorm_fwd"),
    "Gaussian log-likelihood": (["npfk_gauss_nll_fwd", "npfk_gauss_nll_bwd"], "hbm", "gauss_nll_fwd_kernel"),
    "SetConv grid->target": (["npfk_setconv_interp_fwd", "npfk_setconv_interp_bwd"], "hbm", "setconv_interp_fwd_v2_kernel"),
}


def ncu_traffic():
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the hot kernels at the bench's own
    shapes, from the committed `ncu --set full` capture (tools/ncu_targets.py -> profiles/ncu_traffic.json)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)
    return {}


def build_roofline(ksum, total_ms, args, pk, B, Nt, workload=None, steps=None):
    """`roofline` object of the JSON line: the dominant kernel family (largest share of the timed step) against the
    roofline that bounds it, plus the same figures for every other timed family under `families`."""
    workload = workload or args.workload
    steps = steps or args.steps
    all_traffic = ncu_traffic()
    # profiles/ncu_traffic.json: {device kernel: {...}} captured at the C2 shapes, plus {"<workload>": {kernel: {...}}}
    # sections for the captures made at other workloads' shapes
    traffic = all_traffic.get(workload, all_traffic if workload == "attncnp_c2" else {})
    fams = {}
    for fam, (names, bound, dev_kernel) in KERNEL_FAMILIES.items():
        ms = sum(ksum.get(n, {}).get("ms_total", 0.0) for n in names)
        calls = sum(ksum.get(n, {}).get("calls", 0) for n in names)
        if calls == 0:
            continue
        flops = sum(ksum.get(n, {}).get("flops_total", 0.0) for n in names)
        nbytes = sum(ksum.get(n, {}).get("bytes_total", 0.0) for n in names)
        tfs = flops / (ms * 1e-3) / 1e12
        gbs = nbytes / (ms * 1e-3) / 1e9
        e = {"bound": bound, "launches_timed": calls, "avg_launch_ms": ms / calls, "ms_per_step": ms / steps,
             "share_of_step": ms / total_ms, "device_kernel": dev_kernel}
        if bound == "tensor":
            # peak = measured dense bf16 (MEASURED_PEAKS.json); the kernels run kind::tf32, whose dense peak is half
            e.update(achieved=tfs, peak=pk["tflops"], unit="TFLOP/s", frac=tfs / pk["tflops"],
                     frac_of_tf32_peak=tfs / (pk["tflops"] / 2), flops_per_launch=flops / calls)
            if fam.startswith("attention"):
                # head dim 16: a (query, key, head) pair is 64 tensor flop forward / 160 backward but also one ex2 and
                # ~5 / ~8 CUDA-core instructions, so the SM's scalar pipes bound these kernels, not the tensor pipe
                # (DESIGN.md section 3): pairs per clock per SM and the MUFU floor (16 ex2 / clk / SM) say how close they are
                pairs = flops / (160.0 if "backward" in fam else 64.0)
                clk = 1.965e9
                e["per_pair"] = {"pairs_per_launch": pairs / calls,
                                 "pairs_per_clk_per_sm": pairs / (ms * 1e-3) / (148 * clk),
                                 "mufu_floor_ms": pairs / calls / (148 * 16 * clk) * 1e3,
                                 "frac_of_mufu_floor": (pairs / calls / (148 * 16 * clk) * 1e3) / (ms / calls),
                                 "note": "B200: 148 SMs x 16 ex2/clk at 1.965 GHz; FlashAttention-4 at head dim 128 runs at "
                                         "~11 (forward) / ~2.9 (backward) pairs/clk/SM"}
        else:
            e.update(achieved=gbs, peak=pk["hbm_gbs"], unit="GB/s", frac=gbs / pk["hbm_gbs"],
                     algorithmic_bytes_per_launch=nbytes / calls, executed_tflops=tfs)
        t = next((v for k, v in traffic.items() if k.startswith(dev_kernel)), None)
        e["traffic"] = t["bytes_per_launch"] if t else None
        if t:
            e["traffic_note"] = t.get("note", "")
        fams[fam] = e
    if not fams:
        return {"bound": "hbm", "achieved": None, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": None, "traffic": None}
    dom = max(fams, key=lambda f: fams[f]["share_of_step"])
    out = {"kernel": dom}
    out.update(fams[dom])
    out["peak_source"] = pk["source"]
    alg = 3 * FLOPS_FWD_PER_TARGET[workload] * B * Nt / (total_ms / steps * 1e-3) / 1e12
    out["whole_step"] = {"algorithmic_tflops": alg, "frac_of_peak": alg / pk["tflops"]}
    out["families"] = {f: e for f, e in fams.items() if f != dom}
    return out


# ------------------------------------------------------------------------------------------ inference sweep (C5)
def run_infer(args):
    """Forward-only throughput of one resident chunk of tasks (eval mode, no autograd), replicas only: every rank
    predicts its own chunk, nothing is exchanged.  The predictive mean/std stay on the device; the e2e leg copies the
    inputs H2D and reads a per-task summary (mean predictive std, 4 B per task) back."""
    import torch
    import torch.distributed as dist

    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200 import kernels as Kn
    from util_models import build_product, chirp_batch

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (product arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/npf_nccl_%h_%p.log")
        dist.init_process_group("nccl", device_id=dev)
    model_name, B, Nc, Nt, nz = WORKLOADS[args.workload]
    if args.tasks_per_gpu:
        B = args.tasks_per_gpu
    npf.set_mode(args.mode)
    torch.manual_seed(0)
    model = build_product(model_name, 128, nz=nz).to(dev).eval()
    host = {k: v.pin_memory() for k, v in chirp_batch(B, Nc, Nt, 1234 + rank).items() if k != "Y_trgt"}
    static = {k: v.to(dev) for k, v in host.items()}
    h2d_bytes = sum(v.numel() * v.element_size() for v in host.values())
    summary_host = torch.zeros(B).pin_memory()

    def forward():
        with torch.no_grad():
            p = model(static["X_cntxt"], static["Y_cntxt"], static["X_trgt"])[0]
            return p.base_dist.loc, p.base_dist.scale

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        forward()
    timed_names = [n for ns in KERNEL_FAMILIES.values() for n in ns[0]]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    n0 = Kn.launch_count()
    with Kn.KernelTimer(timed_names) as kt:
        total_ms = timed_loop(forward, args.steps)
    launches = Kn.launch_count() - n0
    ksum = kt.summary()
    eager_ms = total_ms / args.steps
    ms_per_step = eager_ms
    step = forward
    use_graph = args.graph in ("on", "auto")
    if use_graph:
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            g_loc, g_scale = forward()
            g_summary = g_scale.mean(dim=(0, 2, 3))
        step = graph.replay
        for _ in range(3):
            step()
        ms_per_step = timed_loop(step, args.steps) / args.steps
    clocks = sampler.stop() if rank == 0 else {}
    value = world * B * Nt / (ms_per_step * 1e-3)

    def step_e2e():
        for k, v in host.items():
            static[k].copy_(v, non_blocking=True)
        if use_graph:
            graph.replay()
            summ = g_summary
        else:
            summ = forward()[1].mean(dim=(0, 2, 3))
        summary_host.copy_(summ, non_blocking=False)

    for _ in range(2):
        step_e2e()
    e2e_ms = timed_loop(step_e2e, args.steps) / args.steps

    # the product-level driver of the sweep (npf_gwwaveform_b200.utils.predict.predict_sharded): this rank's shard of
    # world * B * n_chunks tasks, chunk by chunk from pinned host memory (H2D of the next chunk overlapped with the
    # current forward), per-task summaries kept on the device and copied to the host once at the end; eager launches
    from npf_gwwaveform_b200.utils.predict import predict_sharded

    n_chunks = 4
    host_y = dict(host)

    def source(lo, hi):
        for _ in range(lo, hi, B):
            yield host_y

    def sweep():
        summ, _ = predict_sharded(model, source, n_tasks=world * B * n_chunks, chunk=B, rank=rank, world_size=world)
        return summ.cpu()

    sweep()
    sweep_ms = timed_loop(sweep, 1)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    roofline = build_roofline(ksum, total_ms, args, pk, B, Nt)
    alg = FLOPS_FWD_PER_TARGET[args.workload] * B * Nt / (ms_per_step * 1e-3) / 1e12
    roofline["whole_step"] = {"algorithmic_tflops": alg, "frac_of_peak": alg / pk["tflops"]}
    line = {
        "metric": "target-points/sec forward (batched inference)", "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.mode == "fp32" else "tf32 operands (tcgen05 kind::tf32), f32 accumulate, f32 storage",
        "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "tasks_per_gpu": B, "global_batch": B * world,
                   "n_context": Nc, "n_target": Nt, "mode": args.mode,
                   "step": "forward only (eval mode, no autograd): predictive mean and std of every target stay in HBM",
                   "launch": "CUDA graph replay" if use_graph else "eager launches", "eager_ms_per_step": eager_ms,
                   "parallelism": f"replicas x{world}, no communication",
                   "l2": "activations streamed per step exceed the 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": world * B * Nt / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": h2d_bytes * world, "d2h_bytes_per_step": 4 * B * world,
                "path": "pinned host inputs -> static device inputs -> model.forward (eval) -> per-task mean predictive "
                        "std to host"},
        "sweep": {"api": "npf_gwwaveform_b200.utils.predict.predict_sharded", "tasks": world * B * n_chunks, "chunk": B,
                  "value": world * B * n_chunks * Nt / (sweep_ms * 1e-3), "unit": UNIT, "ms": sweep_ms,
                  "path": "shard_range over ranks -> chunks from pinned host memory, H2D double-buffered on a copy stream -> "
                          "model.forward (eval, eager launches) -> per-task summary on device -> one D2H of the summaries"},
        "gpu_launches": launches, "roofline": roofline,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------ product arm
# strong scaling: the GLOBAL batch BASELINE.json names is fixed and split over the ranks; a rank walks its share in
# micro-batches of at most the per-GPU batch of the weak-scaling run (gradient accumulation inside the captured step)
GLOBAL_BATCH = {"attncnp_c2": 256, "convcnp_c3": 256, "attnlnp_c4": 128, "cnp_c1": 64}


def measure_training(args, workload, dev, rank, world, scaling, steps, warmup, with_timer=True, with_e2e=True):
    """Build the model of `workload`, run `warmup` + `steps` optimisation steps (eager, timed per kernel), capture the
    step into CUDA graphs and time `steps` replays.  Returns a dict of raw measurements (rank-0 view; the timings are
    the max over ranks)."""
    import torch
    import torch.distributed as dist

    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200 import kernels as Kn
    from npf_gwwaveform_b200.training import FlatAdamTrainer, shard_range
    from util_models import build_product, chirp_batch

    model_name, B, Nc, Nt, nz = WORKLOADS[workload]
    if args.tasks_per_gpu and workload == args.workload:
        B = args.tasks_per_gpu
    micro = 1
    if scaling == "strong":
        G = GLOBAL_BATCH[workload]
        assert G % world == 0, f"global batch {G} does not split over {world} ranks"
        per_rank = G // world
        micro = max(1, (per_rank + B - 1) // B)
        assert per_rank % micro == 0
        B = per_rank // micro
    torch.manual_seed(0)                            # identical initial weights on every rank
    model = build_product(model_name, 128, nz=nz).to(dev).train()
    crit = (npf.ELBOLossLNPF if "LNP" in model_name else npf.CNPFLoss)().train()
    trainer = FlatAdamTrainer(model, crit, lr=1e-3, n_buckets=int(os.environ.get("NPFK_BUCKETS", "2")))

    # synthetic shard resident in HBM for `value`, pinned on the host for `e2e`.  weak: every rank draws its own tasks
    # (seed 1234 + rank); strong: rank r takes its slice of ONE global batch (seed 1234), micro-batch by micro-batch
    if scaling == "strong":
        whole = chirp_batch(GLOBAL_BATCH[workload], Nc, Nt, 1234)
        lo, hi = shard_range(GLOBAL_BATCH[workload], rank, world)
        hosts = [{k: v[lo + i * B:lo + (i + 1) * B].contiguous().pin_memory() for k, v in whole.items()} for i in range(micro)]
        del whole
    else:
        hosts = [{k: v.pin_memory() for k, v in chirp_batch(B, Nc, Nt, 1234 + rank).items()}]
    residents = [{k: v.to(dev, non_blocking=True) for k, v in h.items()} for h in hosts]
    h2d_bytes = sum(v.numel() * v.element_size() for h in hosts for v in h.values())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(step_fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            step_fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_eager():        # micro > 1: eager warm-up steps treat every micro-batch as a step of its own
        loss = None
        for r in residents:
            loss = trainer.step(r)
        return loss

    for _ in range(max(warmup, 3)):
        step_eager()
    timed_names = [n for ns in KERNEL_FAMILIES.values() for n in ns[0]] if with_timer else []
    n0 = Kn.launch_count()
    with Kn.KernelTimer(timed_names) as kt:
        total_ms = timed_loop(step_eager, steps)
    launches = Kn.launch_count() - n0
    res = dict(workload=workload, model=model_name, tasks_per_gpu=B * micro, micro_batches=micro, micro_batch=B, Nc=Nc, Nt=Nt,
               nz=nz, ksum=kt.summary(), eager_total_ms=total_ms, eager_ms_per_step=total_ms / steps, launches=launches,
               h2d_bytes=h2d_bytes, graph=False)
    ms_per_step = res["eager_ms_per_step"]
    step_resident = step_eager
    loss_host = torch.zeros(1).pin_memory()

    def step_e2e_eager():
        loss = None
        for h in hosts:
            batch = {k: v.to(dev, non_blocking=True) for k, v in h.items()}      # H2D of this step's inputs
            loss = trainer.step(batch)
        loss_host.copy_(loss.reshape(1), non_blocking=False)                        # D2H read of the step's result
        return loss_host

    step_e2e = step_e2e_eager
    if args.graph in ("on", "auto"):
        # the same step (zero + forward + loss + backward + bucketed NCCL all-reduce + Adam) replayed as CUDA graphs;
        # the kernels and their order are those of the eager loop above, which also gave the per-kernel event times
        trainer.capture(residents[0], micro_batches=micro)

        def step_graph():
            loss = None
            for i, r in enumerate(residents):
                loss = trainer.step_graphed(r, first=i == 0, last=i == micro - 1)
            return loss

        def step_e2e_graph():
            # input pipeline: the H2D copy of the next (micro-)batch was started while the previous one computed; this
            # one waits for it, copies it D2D into the graph's static inputs, replays, and starts the next H2D before
            # the loss is read back: every input byte crosses PCIe once per step inside the timed region
            loss = None
            for i, h in enumerate(hosts):
                loss = trainer.step_graphed(h, first=i == 0, last=i == micro - 1)
                trainer.prefetch(hosts[(i + 1) % micro])
            loss_host.copy_(loss.reshape(1), non_blocking=False)
            return loss_host

        for _ in range(3):
            step_graph()
        ms_per_step = timed_loop(step_graph, steps) / steps
        step_e2e, step_resident = step_e2e_graph, step_graph
        res["graph"] = True
    res["ms_per_step"] = ms_per_step
    res["value"] = world * B * micro * Nt / (ms_per_step * 1e-3)
    if with_e2e:
        for _ in range(2):
            step_e2e()
        e2e_ms = timed_loop(step_e2e, steps) / steps
        res["e2e_ms"] = e2e_ms
        res["e2e_value"] = world * B * micro * Nt / (e2e_ms * 1e-3)
    res["final_loss"] = float(step_resident().item())
    res["trainer"] = trainer
    res["residents"] = residents
    return res


def dp_check(res, dev, rank, world):
    """NCCL correctness of the data-parallel step: the all-reduced mean gradient of the ranks' shards must equal the
    gradient rank 0 gets by walking ALL shards itself on the same weights (shards gathered over NCCL, no optimiser
    step).  BatchNorm models use per-replica statistics (reference semantics on a shard), which this reproduces shard by
    shard.  Returns the relative L2 difference (rank 0) or None."""
    import torch
    import torch.distributed as dist

    trainer = res["trainer"]
    model, crit = trainer.model, trainer.criterion
    shard = res["residents"][0]
    was = {n: b.clone() for n, b in model.named_buffers()}      # BatchNorm running statistics must not drift

    def grad_of(batch):
        trainer.flat_grad.zero_()
        crit(model(**batch), batch["Y_trgt"]).backward()
        return trainer.flat_grad.clone()

    # the trainer's post-accumulate hooks would launch bucket all-reduces from inside these backward passes (and rank 0
    # runs more of them than the other ranks): route the hook events to a no-op for the duration of the check
    trainer._event_cb = lambda _i: None

    g_local = grad_of(shard)
    g_dp = g_local.clone()
    dist.all_reduce(g_dp, op=dist.ReduceOp.SUM)
    g_dp /= world
    gathered = {k: [torch.empty_like(v) for _ in range(world)] for k, v in shard.items()}
    for k, v in shard.items():
        dist.all_gather(gathered[k], v.contiguous())
    out = None
    if rank == 0:
        g_ref = torch.zeros_like(g_local)
        for r in range(world):
            g_ref += grad_of({k: gathered[k][r] for k in shard})
        g_ref /= world
        out = float((g_dp - g_ref).norm() / g_ref.norm().clamp_min(1e-30))
    trainer._event_cb = None
    trainer.flat_grad.zero_()
    with torch.no_grad():
        for n, b in model.named_buffers():
            b.copy_(was[n])
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.neuralproc import set_lazy_validation

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (product arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner (NCCL_DEBUG=VERSION/WARN/INFO) on stdout: keep stdout to the ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/npf_nccl_%h_%p.log")
        # a mismatched collective must fail fast, not sit in NCCL's default 10-minute watchdog window
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    npf.set_mode(args.mode)
    set_lazy_validation(True)                       # range flag still computed every step, read one step late

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    res = measure_training(args, args.workload, dev, rank, world, args.scaling, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else {}
    check = dp_check(res, dev, rank, world) if world > 1 else None
    res.pop("trainer").close()
    res.pop("residents")
    torch.cuda.empty_cache()

    # secondary block: the other training configurations BASELINE.json names, measured in the same process so that the
    # driver's 1/2/4/8-GPU run covers them: configs[2] ConvCNP-UNet at 32 tasks per GPU (= batch 256 over 8 GPUs, weak
    # scaling) and configs[3] AttnLNP with the global batch fixed at 128 (strong scaling, micro-batches of 16 tasks)
    secondary = {}
    if args.workload == "attncnp_c2" and not args.no_secondary:
        for wl, sc in (("convcnp_c3", "weak"), ("attnlnp_c4", "strong")):
            r2 = measure_training(args, wl, dev, rank, world, sc, max(3, args.steps // 2), 3, with_timer=True, with_e2e=False)
            r2.pop("trainer").close()
            r2.pop("residents")
            torch.cuda.empty_cache()
            secondary[wl] = (r2, sc)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    B, Nc, Nt, nz = res["tasks_per_gpu"], res["Nc"], res["Nt"], res["nz"]
    roofline = build_roofline(res["ksum"], res["eager_total_ms"], args, pk, B, Nt, args.workload)
    use_graph = res["graph"]
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32" if args.mode == "fp32" else "tf32 operands (tcgen05 kind::tf32), f32 accumulate, f32 storage",
        "data": "synthetic",
        "config": {"workload": WORKLOAD_TEXT[args.workload], "tasks_per_gpu": B, "global_batch": B * world,
                   "micro_batches_per_step": res["micro_batches"],
                   "n_context": Nc, "n_target": Nt, "n_z": nz, "mode": args.mode,
                   "step": "forward + loss + backward + gradient all-reduce + Adam",
                   "launch": ("the whole step (zero grads, forward, loss, backward, bucketed NCCL all-reduce forked off in the "
                              "middle of backward, Adam) replayed as one CUDA graph" if use_graph else "eager launches"),
                   "eager_ms_per_step": res["eager_ms_per_step"],
                   "parallelism": f"data-parallel over tasks x{world}",
                   "l2": "activations streamed per step (several GB) exceed the 126 MB L2; no explicit flush"},
        "clocks": clocks,
        "e2e": {"value": res["e2e_value"], "unit": UNIT, "ms_per_step": res["e2e_ms"],
                "h2d_bytes_per_step": res["h2d_bytes"] * world, "d2h_bytes_per_step": 4 * world,
                "path": ("pinned host tensors -> H2D on a copy stream into a staging set while the previous step runs "
                         "(FlatAdamTrainer.prefetch) -> D2D into the graph's static inputs -> graph replay (forward, "
                         "loss, backward, all-reduce, Adam) -> loss to host (blocking D2H every step)") if use_graph
                else "pinned host tensors -> .to(device) -> model(**batch) -> criterion -> backward -> Adam -> loss to host"},
        "gpu_launches": res["launches"],
        "roofline": roofline,
        "final_loss": res["final_loss"],
    }
    if check is not None:
        line["dp_check"] = {"what": "||allreduce(mean of shard gradients) - gradient of all shards on rank 0|| / ||.||",
                            "rel_l2": check, "ok": bool(check < 5e-3)}
        assert check < 5e-3, f"data-parallel gradient differs from the single-rank gradient: {check}"
    for wl, (r2, sc) in secondary.items():
        secondary[wl] = {
            "workload": WORKLOAD_TEXT[wl], "scaling": sc, "value": r2["value"], "unit": UNIT, "ms_per_step": r2["ms_per_step"],
            "tasks_per_gpu": r2["tasks_per_gpu"], "global_batch": r2["tasks_per_gpu"] * world,
            "micro_batches_per_step": r2["micro_batches"], "eager_ms_per_step": r2["eager_ms_per_step"],
            "gpu_launches": r2["launches"], "final_loss": r2["final_loss"],
            "roofline": build_roofline(r2["ksum"], r2["eager_total_ms"], args, pk, r2["tasks_per_gpu"], r2["Nt"], wl,
                                       steps=max(3, args.steps // 2)),
        }
    if secondary:
        line["secondary"] = secondary
    if world == 1 and not args.no_cpu_baseline:
        rate, desc, cores, _ = cpu_reference_rate(args.workload, seconds=args.cpu_seconds)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    """the ONE JSON line, on the process's original stdout"""
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    args = parse()
    # Libraries print on stdout behind our back (NCCL's version banner under NCCL_DEBUG=VERSION is a plain printf):
    # keep a private handle on the real stdout for the JSON line and point fd 1 at stderr for everything else.
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    elif args.workload.endswith("_infer"):
        run_infer(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
