"""Headline benchmark (BASELINE.json): frames/s of VDA-L 518x518 32-frame windows on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--arch vitl]

A "step" is one pass of the hot path (VideoDepthAnything.forward, video_depth.py:61-68) over one
synthetic 32-frame window per rank.  N>1 is launched by torchrun, one rank per GPU; windows are
independent so ranks share nothing on the data path ("scaling": "weak").
  value : 32*K*N / max-over-ranks device time, the window already resident in HBM.
  e2e   : the same through the public module call with HOST buffers: pinned fp32 window -> H2D ->
          forward -> D2H of the fp32 depth, all inside the timed region.
  roofline : the tcgen05 GEMM kernel (gemm_tc_kernel, A_PLAIN launches) timed per launch with CUDA
          events on the launching stream over a second pass of the same K steps launched eagerly
          (the timed pass replays a CUDA graph, which cannot hold events); FLOPs = 2*M*N*K per launch.
  cpu_baseline : the oracle (CPU restatement of the reference forward) on the host cores, on a
          bounded sample (a T=--cpu-frames window of the same architecture and resolution).
--impl reference runs only that CPU arm (rank 0) with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frames/sec VDA-L 518x518 32-frame windows"
UNIT = "frames/s"
NCU_FC1 = "r01c_ncu_fc1.json"   # ncu --set full capture of the fc1 launch the roofline's `traffic` is read from


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--arch", default="vitl", choices=["vits", "vitb", "vitl"])
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--size", type=int, default=518)
    ap.add_argument("--cpu-frames", type=int, default=2, help="frames of the CPU-baseline sample window")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--torch-eager-gpu", action="store_true",
                    help="also time the oracle (plain torch ops, the reference's own op sequence) on this GPU: fp32 and "
                         "autocast bf16 -- SURVEY.md 8(d)'s second comparand; reported, never the measured product")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="fp32 = the FP32-accuracy mode (split-bf16 products, DESIGN.md 4b); not the headline config")
    return ap.parse_args()


def config_of(a):
    prec = "BF16" if a.precision == "bf16" else "FP32-accuracy mode (3-term split-bf16 products)"
    return {"workload": f"Video-Depth-Anything ({a.arch}) {prec}, one {a.frames}-frame window {a.size}x{a.size} per GPU "
                        f"per step (BASELINE configs[1])",
            "arch": a.arch, "frames_per_window": a.frames, "height": a.size, "width": a.size,
            "weights": "random-init (seeded)", "sharding": "one independent window per rank per step",
            "l2": "per-step working set (several GB of activations) is larger than the 126 MB L2; no explicit flush"}


# ----------------------------------------------------------------------------- CPU arm (oracle)
def cpu_arm(a, steps, warmup):
    """Times the oracle forward (fp32, torch CPU, all host threads) on a T=cpu_frames window."""
    import torch

    from oracle import refload
    from oracle import vda_oracle as O

    torch.set_grad_enabled(False)
    cores = torch.get_num_threads()
    sd = refload.seeded_state_dict(a.arch, 0)
    T = a.cpu_frames
    x = torch.randn(1, T, 3, a.size, a.size, generator=torch.Generator().manual_seed(0))
    for _ in range(warmup):
        O.forward(sd, x, a.arch)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.forward(sd, x, a.arch)
    dt = (time.perf_counter() - t0) / steps
    return {"value": T / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle forward on a {T}-frame {a.size}x{a.size} {a.arch} window, fp32 torch CPU, "
                      f"{steps} timed pass(es) of {dt:.2f} s"}, dt


def torch_eager_gpu_arm(a, steps=3, warmup=2):
    """The oracle's op sequence (what the reference executes: F.linear / naive softmax attention / conv2d /
    interpolate ...) as eager PyTorch on the same B200, full window, fp32 and autocast(bf16)."""
    import torch

    from oracle import refload
    from oracle import vda_oracle as O

    torch.set_grad_enabled(False)
    sd = {k: v.cuda() for k, v in refload.seeded_state_dict(a.arch, 0).items()}
    x = torch.randn(1, a.frames, 3, a.size, a.size, generator=torch.Generator().manual_seed(0)).cuda()
    out = {}
    for name, ctx in (("fp32", torch.autocast("cuda", enabled=False)),
                      ("autocast_bf16", torch.autocast("cuda", dtype=torch.bfloat16)),
                      ("autocast_fp16", torch.autocast("cuda", dtype=torch.float16))):
        with ctx:
            for _ in range(warmup):
                O.forward(sd, x, a.arch)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(steps):
                O.forward(sd, x, a.arch)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out[name] = {"frames_per_s": a.frames / ms * 1e3, "ms_per_window": ms}
    out["note"] = (f"oracle forward (eager torch {torch.__version__}, naive attention as attention.py:53-59, TF32 off), "
                   f"{a.frames}-frame {a.size}x{a.size} {a.arch} window, {steps} timed passes")
    return out


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(a.steps, 3)), 1 if a.warmup > 0 else 0
    cb, dt = cpu_arm(a, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_of(a),
            "cpu_baseline": cb, "gpu_launches": 0,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- clocks sampler
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples, self.stop_flag, self.index = [], False, index
        self.thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        try:
            p = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                  "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
        except OSError:
            return
        self.proc = p
        for line in p.stdout:
            if self.stop_flag:
                break
            self.samples.append([s.strip() for s in line.split(",")])
        p.kill()

    def start(self):
        self.thread.start()

    def stop(self):
        self.stop_flag = True
        time.sleep(0.15)
        if hasattr(self, "proc"):
            self.proc.kill()
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = max([int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples if len(s) >= 6 for n, v in zip(names, s[2:6]) if v.lower() == "active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons,
                "samples": len(sm)}


# ----------------------------------------------------------------------------- CUDA arm
def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return p["bf16_tflops_sustained"], p["hbm_gbs"], "measured (MEASURED_PEAKS.json, sustained bf16)"
    except Exception:
        return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes (read + write) of the largest plain-GEMM launch (vitl fc1) from the committed
    `ncu --set full` capture under profiles/ -- a profiler artefact, not measured in this run."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", NCU_FC1)))

        def mb(k):
            v = d[k]
            if isinstance(v, dict):
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[v["unit"]]
                return float(v["value"]) * scale
            num, unit = v.split()
            return float(num) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]

        alg = sum(d["algorithmic_bytes"].values())
        return mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum"), alg
    except Exception:
        return None, None


def run_ours(a):
    import torch
    import torch.distributed as dist

    import vda_b200.lib as L
    from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_
    from video_depth_anything.video_depth import VideoDepthAnything

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    torch.set_grad_enabled(False)
    torch.manual_seed(0)
    model = randomize_zero_init_(VideoDepthAnything(**MODEL_CONFIGS[a.arch])).to(dev).eval()
    T, H, W = a.frames, a.size, a.size
    g = torch.Generator().manual_seed(100 + rank)
    x_host = torch.randn(1, T, 3, H, W, generator=g).pin_memory()
    x_dev = x_host.to(dev)
    fp32 = a.precision == "fp32"
    eng = model.engine(fp32)

    # ---- device-resident timing (value): K window forwards, each one CUDA-graph replay of the
    # window's ~270 kernel launches
    for _ in range(a.warmup):
        eng.forward(x_dev[0])
    barrier()
    clocks = Clocks(local) if rank == 0 else None
    if clocks:
        clocks.start()
    n0 = L.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        eng.forward(x_dev[0])
    e1.record()
    barrier()
    launches = L.LAUNCHES - n0
    ms = max_over_ranks(e0.elapsed_time(e1))
    d_last = eng.forward(x_dev[0])
    depth_stats = [float(d_last.min()), float(d_last.mean()), float(d_last.max())]
    clk = clocks.stop() if clocks else None

    # ---- the same K steps launched eagerly with CUDA events around every tensor-core launch
    # (events cannot be recorded inside a graph replay): per-kernel durations for the roofline
    L.PROFILE = []
    e0.record()
    for _ in range(a.steps):
        eng.forward(x_dev[0])
    e1.record()
    barrier()
    prof, L.PROFILE = L.PROFILE, None
    ms_eager = e0.elapsed_time(e1)

    # ---- end to end through the public API with HOST buffers: every step copies its own pinned fp32
    # window to the device and its fp32 depth back to pinned host memory, all inside the timed region
    # (forward_pipelined overlaps window i+1's H2D and window i-1's D2H with window i's kernels)
    outs_host = [torch.empty(T, H, W, dtype=torch.float32).pin_memory() for _ in range(2)]

    def e2e_run(k):
        return model.forward_pipelined((x_host[0] for _ in range(k)), out=[outs_host[i & 1] for i in range(k)], fp32=fp32)

    e2e_run(2)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    e2e_run(a.steps)          # returns after the last D2H has completed
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), 0.0))
    ms_e2e_wall = (time.perf_counter() - t0) * 1e3

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    kinds = {}
    for kind, flops, s, e in prof:
        k = kinds.setdefault(kind, [0, 0.0, 0.0])
        k[0] += 1
        k[1] += flops
        k[2] += s.elapsed_time(e)
    tf_peak, hbm_peak, peak_src = peaks()
    hbm = {k[4:]: v for k, v in kinds.items() if k.startswith("hbm:")}
    kinds = {k: v for k, v in kinds.items() if not k.startswith("hbm:")}
    n_g, f_g, t_g = kinds.get("gemm", [1, 0.0, 1e-9])
    achieved = f_g / (t_g * 1e-3) / 1e12
    total_frames = T * a.steps * world
    line = {
        "metric": METRIC, "value": total_frames / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if not fp32 else "bf16x3 (fp32-accurate)", "data": "synthetic",
        "config": config_of(a),
        "p50_window_latency_ms": ms / a.steps, "ms_per_step_eager_instrumented": ms_eager / a.steps,
        "e2e": {"value": total_frames / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": x_host.numel() * 4, "d2h_bytes_per_step": T * H * W * 4,
                "ms_per_step": ms_e2e / a.steps, "wall_ms_per_step": ms_e2e_wall / a.steps,
                "api": "VideoDepthAnything.forward_pipelined (pinned host in/out, copies on side streams)"},
        "gpu_launches": launches, "depth_min_mean_max": depth_stats,
        "clocks": clk,
        "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (plain GEMM launches: encoder/motion Linears, 1x1 convs)",
                     "achieved": achieved, "peak": tf_peak, "unit": "TFLOP/s", "frac": achieved / tf_peak,
                     "traffic": ncu_traffic()[0], "traffic_algorithmic_bytes": ncu_traffic()[1],
                     "traffic_note": "DRAM read+write bytes of one fc1 launch (M=43840,N=4096,K=1024), ncu --set full, "
                                     f"profiles/{NCU_FC1}",
                     "peak_source": peak_src, "launches": n_g,
                     "flops_per_launch": f_g / n_g, "ms_per_launch": t_g / n_g},
        "breakdown_ms_per_step": {k: v[2] / a.steps for k, v in kinds.items()},
        "breakdown_tflops": {k: v[1] / (v[2] * 1e-3) / 1e12 for k, v in kinds.items() if v[2] > 0},
        # bandwidth-bound kernels: algorithmic bytes / event time, against the measured HBM peak
        "hbm_kernels": {k: {"ms_per_step": v[2] / a.steps, "GBps": v[1] / (v[2] * 1e-3) / 1e9,
                            "frac_of_hbm_peak": v[1] / (v[2] * 1e-3) / 1e9 / hbm_peak} for k, v in hbm.items() if v[2] > 0},
    }
    if world == 1 and not a.no_cpu_baseline:
        line["cpu_baseline"], _ = cpu_arm(a, 1, 1)
    if world == 1 and a.torch_eager_gpu:
        del model
        torch.cuda.empty_cache()
        line["torch_eager_same_gpu"] = torch_eager_gpu_arm(a)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
