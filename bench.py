"""Headline benchmark: spectral-ViT training images/s on synthetic data + mixer-kernel roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c4|c2|c3|c5|c1] [--batch B] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One JSON line on rank 0.  "step" = one optimisation step (forward, loss, backward with the fused
filter-gradient reduction, DDP all-reduce, fused AdamW) of the configured Spectre-ViT on one batch per
GPU.  Default workload: BASELINE.json configs[3] (ViT-S/16 FFT-mix, 224^2 -> 14x14 grid, bf16, batch 2048 per GPU =
configs/vit_s16_fft.py:bench_batch_size; --batch 256 gives SURVEY.md's batch) -- the configuration the north-star's
scaling target is quoted on.

  value     images/s, inputs resident in HBM, CUDA events, max over ranks.  The step is captured once in a CUDA graph
            (train.Trainer(cuda_graph=True)) and the timed region replays it; --no-cuda-graph times eager steps
  e2e       same through the public API with HOST buffers: every step copies the batch from pinned host
            memory and reads the loss back
  roofline  dominant mixer kernel (the fused backward), timed with CUDA events on the launching stream: inside the
            timed steps with --no-cuda-graph; by default (events cannot be recorded inside a replayed graph) in an
            eager pass of the same steps right after the timed replays; algorithmic bytes = 3*B*T*E*s + 2*F / mean launch time
  cpu_baseline  the oracle port of the reference's PyTorch path (reference model classes' mirror with the
            torch.fft / Haar composition in the mix_layer slot) on the host cores, bounded sample
--impl reference runs only that CPU arm (rank 0; other ranks exit 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "spectral-ViT train images/sec (synthetic); mixer-kernel HBM GB/s vs peak"
WORKLOADS = {
    "c1": "Spectre-ViT tiny FFT-mix, CIFAR-100-shaped 32x32, patch 4 (8x8 grid), fp32",
    "c2": "Spectre-ViT tiny FFT-mix, MNIST-shaped 28x28, patch 4 (7x7 grid), fp32",
    "c3": "Spectre-ViT DWT-mix (Haar) CIFAR-100-shaped, bf16",
    "c4": "ViT-S/16 FFT-mix 224^2 (14x14 grid) synthetic ImageNet-shaped, bf16",
    "c5": "ViT-B/16 spectral student distilled from random-init DINOv3 ViT-B/16 teacher, 224^2, bf16",
}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def oracle_factory(kind):
    """mix_layer occupant = the oracle's torch.fft / Haar composition (CPU baseline only)."""
    from oracle.mixers_ref import OracleDWTMix, OracleFFTMix

    cls = OracleFFTMix if kind == "fft" else OracleDWTMix
    return lambda d_model, seq, nhead, out: cls(d_model, seq, nhead, out)


REF_DIR = os.path.join(ROOT, "baseline", "_ref")   # the unmodified reference, installed by baseline/install_ref.sh


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_DIR, "spectre_vit")) and os.path.isdir(os.path.join(REF_DIR, "dinov3"))


def reference_teacher(cfg):
    """Config 5's teacher as the reference builds it (spectre_vit/repl/train.py:60-83, distillation.py:5-22):
    ``DinoClassifier(dinov3_vitb16(pretrained=False), num_classes, embed_dim=768)``, random init, frozen."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    from dinov3.hub.backbones import dinov3_vitb16
    from spectre_vit.distillation import DinoClassifier

    return DinoClassifier(backbone=dinov3_vitb16(pretrained=False), num_classes=cfg.num_classes,
                          embed_dim=cfg.teacher_embed_dim)


def reference_model(cfg):
    """The reference's OWN ``SpectreViT`` (spectre_vit/models/spectre/spectre.py:159-202) with the oracle's torch.fft /
    Haar composition placed in every layer's ``mix_layer`` slot (the reference has no module of its own there)."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    from spectre_vit.models.spectre.spectre import SpectreViT as RefViT

    m = RefViT(img_size=cfg.img_size, patch_size=cfg.patch_size, in_channels=cfg.in_channels, num_classes=cfg.num_classes,
               embed_dim=cfg.embed_dim, num_encoders=cfg.num_encoders, num_heads=cfg.num_heads, hidden_dim=cfg.hidden_dim,
               dropout=cfg.dropout, activation=cfg.activation)
    make = oracle_factory(cfg.mixer)
    T = (cfg.img_size // cfg.patch_size) ** 2 + 1
    for layer in m.encoder_blocks.layers:
        layer.mix_layer = make(cfg.embed_dim, T, cfg.num_heads, cfg.embed_dim)
    return m


def cpu_reference_leg(cfg, steps, warmup, batch, label):
    """The reference's PyTorch path on the host cores: fwd + bwd + AdamW, fp32.  With baseline/_ref present the model
    (and config 5's teacher) are the reference's own classes (kind "reference"); otherwise the repo's mirror of them
    (kind "port", pinned to the reference by tests/golden)."""
    from vit_spectre_experiments_b200.train import Trainer

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg_cpu = type(cfg)(**{**vars(cfg), "precision": "fp32"})
    kind = "port"
    model = teacher = None
    if reference_available():
        torch.manual_seed(cfg.random_seed)
        model, kind = reference_model(cfg_cpu), "reference"
        if getattr(cfg, "distill", False):
            teacher = reference_teacher(cfg_cpu)
    tr = Trainer(cfg_cpu, "cpu", mixer=oracle_factory(cfg.mixer), ddp=False, model=model, teacher=teacher)
    img, lab = tr.synthetic_batch(batch)
    for _ in range(warmup):
        tr.step(img, lab)
    t0 = time.perf_counter()
    for _ in range(steps):
        tr.step(img, lab)
    dt = time.perf_counter() - t0
    what = ("reference SpectreViT class + oracle mixer in its slot" if kind == "reference"
            else "mirror of the reference classes + oracle mixer")
    return {"value": round(batch * steps / dt, 2), "unit": "images/s", "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{label}: {what}; {steps} steps x {batch} images fwd+bwd+AdamW on CPU fp32 (torch.fft has no "
                      f"bf16), {warmup} warm-up, {dt:.1f} s"}, dt / steps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-batch", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false",
                    help="eager steps with DDP buckets instead of the default: the whole step (fwd, bwd, gradient "
                         "all-reduce, AdamW) captured in one CUDA graph and replayed")
    args = ap.parse_args()

    from vit_spectre_experiments_b200.configs import load_named
    from vit_spectre_experiments_b200.train import Trainer, dist_env, init_distributed

    cfg = load_named(args.config)
    rank, local_rank, world = dist_env()
    batch = args.batch or getattr(cfg, "bench_batch_size", cfg.batch_size)
    cpu_batch = args.cpu_batch or {"c4": 64, "c5": 16, "c3": 256, "c1": 256, "c2": 2048}[args.config]   # ~10-20 s of CPU work
    grid = cfg.img_size // cfg.patch_size
    T, E, L = grid * grid + 1, cfg.embed_dim, cfg.num_encoders
    config = {"workload": WORKLOADS[args.config], "name": args.config, "mixer": cfg.mixer, "grid": f"{grid}x{grid}",
              "tokens": T, "embed_dim": E, "layers": L, "per_gpu_batch": batch, "global_batch": batch * world,
              "parallelism": f"dp{world}", "optimizer": "AdamW(fused)", "precision": cfg.precision,
              "l2": "per-step activation working set >> 126 MB L2; no flush needed"}

    if args.impl == "reference":
        if rank != 0:
            return
        steps = max(1, min(args.steps, 3))
        warm = max(1, min(args.warmup, 1))
        cb, ms = cpu_reference_leg(cfg, steps, warm, cpu_batch, args.config)
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "images/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": round(ms, 2), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "config": {**config, "per_gpu_batch": cpu_batch,
                                                                                 "global_batch": cpu_batch},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the spectral mixers have no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    rank, local_rank, world = init_distributed()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    torch.backends.cuda.matmul.allow_tf32 = True     # reference: train.py:36-39
    torch.backends.cudnn.allow_tf32 = True

    from vit_spectre_experiments_b200 import _lib
    from vit_spectre_experiments_b200.functional import TIMER

    _lib.load()
    teacher = None
    if getattr(cfg, "distill", False):
        if reference_available():
            teacher = reference_teacher(cfg)
            config["teacher"] = "reference DinoClassifier(dinov3_vitb16(pretrained=False)) from baseline/_ref, frozen"
        else:
            config["teacher"] = "stock ViT-B/16 of the same sizes (baseline/_ref absent: run baseline/install_ref.sh)"
    tr = Trainer(cfg, dev, teacher=teacher, cuda_graph=args.cuda_graph)
    config["cuda_graph"] = bool(args.cuda_graph)
    img, lab = tr.synthetic_batch(batch, rank)
    warm = max(3, args.warmup)
    K = args.steps

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    # ---------------- device-resident leg
    for _ in range(warm):
        tr.step(img, lab)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    TIMER.reset()
    TIMER.enabled = not args.cuda_graph      # events cannot be recorded inside a replayed graph: see below
    launches0 = _lib.launch_counter["kernels"]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(K):
        loss = tr.step(img, lab)
    t1.record()
    barrier()
    TIMER.enabled = False
    ms_total = t0.elapsed_time(t1)
    launches = _lib.launch_counter["kernels"] - launches0
    clocks = sampler.stop() if sampler else None
    if args.cuda_graph:
        # per-kernel timings for the roofline block: the same step run eagerly (same kernels, same shapes) right after
        # the timed graph replays, bracketed by CUDA events on the launching stream
        TIMER.enabled = True
        for _ in range(min(K, 5)):
            tr._eager_step(img, lab)
        TIMER.enabled = False
        config["roofline_timing"] = "eager pass after the timed graph replays"
    ksum = TIMER.summary()
    final_loss = float(loss.item())

    # ---------------- end-to-end leg: pinned host batch -> H2D every step, loss read back every step.
    # The public feed is train.HostBatchPrefetcher: the copy of batch i+1 is enqueued on a side stream before
    # step i runs, so each timed step still copies one whole batch (K copies in the timed region) and syncs on
    # its loss, but PCIe time overlaps compute.
    from vit_spectre_experiments_b200.train import HostBatchPrefetcher

    himg, hlab = tr.synthetic_batch(batch, rank, pinned_host=True)
    feed = HostBatchPrefetcher(dev, img, lab)
    feed.stage(himg, hlab)

    def e2e_step():
        dimg, dlab = feed.next()
        feed.stage(himg, hlab)            # next step's batch: H2D overlaps this step's compute
        loss = tr.step(dimg, dlab)
        feed.release()
        return float(loss.item())

    for _ in range(3):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)

    t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = t.tolist()

    if rank == 0:
        peaks, peak_src = measured_peaks()
        kind = cfg.mixer
        key = f"{kind}mix_bwd"
        roof = None
        kernels = {}
        s = 2 if cfg.precision == "bf16" else 4
        F = (grid * (grid // 2 + 1) * E * 8) if kind == "fft" else (grid * grid * E * 4)
        traffic = {}
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic = json.load(f)
        except Exception:
            pass
        # mean_ms = the main kernel of the call; mean_call_ms = the whole C-ABI call (a backward mixer call ends
        # with a few-microsecond reduction kernel of the per-slab parameter gradients)
        for name, (n, call_ms, meta, mean_ms) in ksum.items():
            if len(meta) != 5:      # row-wise tails (norm_act_*:N:Ns:act:in1): rows, N, element size, algorithmic bytes
                nb = meta[3]
                kernels[name] = {"launches": n, "mean_ms": round(mean_ms, 4), "algo_bytes": nb,
                                 "GBps": round(nb / mean_ms / 1e6, 1), "frac": round(nb / mean_ms / 1e6 / peaks["hbm_gbs"], 4)}
                continue
            Bk, Hk, Wk, Ek, sz = meta
            nbytes = (3 if name.endswith("bwd") else 2) * Bk * T * Ek * sz + (2 if name.endswith("bwd") else 1) * F
            kernels[name] = {"launches": n, "mean_ms": round(mean_ms, 4), "mean_call_ms": round(call_ms, 4),
                             "algo_bytes": nbytes, "GBps": round(nbytes / mean_ms / 1e6, 1),
                             "GBps_call": round(nbytes / call_ms / 1e6, 1), "act_bytes_per_elem": sz}
        if key in kernels:
            k = kernels[key]
            roof = {"bound": "hbm", "kernel": key, "achieved": k["GBps"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": round(k["GBps"] / peaks["hbm_gbs"], 4), "peak_source": f"{peak_src} (MEASURED_PEAKS.json)",
                    "algo_bytes_per_launch": k["algo_bytes"], "mean_launch_ms": k["mean_ms"],
                    "mean_call_ms": k["mean_call_ms"], "frac_whole_call": round(k["GBps_call"] / peaks["hbm_gbs"], 4),
                    "traffic": traffic.get(f"{args.config}:{key}"),
                    "traffic_source": "profiles/ncu_traffic.json (one ncu --set full capture of this kernel at this shape; "
                                      "not re-measured in this run)"}
        # ---- model-level roofline: t_roof/step = memory-bound bytes / HBM peak + GEMM flops /
        # sustained tensor peak.  GEMM flops: the two Linears of every encoder layer + patch projection + head, x3 for
        # forward + both backward GEMMs.  Memory-bound bytes: the algorithmic bytes of every own kernel launched in a
        # step (mixers + row-wise tails, counted from the calls above) + the image read + AdamW's 28 B per parameter.
        nparam = sum(p.numel() for p in tr.model.parameters())
        hid, P, Cin = cfg.hidden_dim, cfg.patch_size, cfg.in_channels
        gemm_flops_img = 3 * 2 * (L * 2 * T * E * hid + (T - 1) * Cin * P * P * E + E * cfg.num_classes)
        if getattr(cfg, "distill", False):   # frozen teacher forward (config 5): 12 E^2 MACs per token and layer + attention
            Et, Lt, Tt = cfg.teacher_embed_dim, cfg.teacher_depth, T + 4     # DINOv3 ViT-B/16: cls + 4 storage tokens
            gemm_flops_img += 2 * Lt * (12 * Et * Et * Tt + 2 * Tt * Tt * Et) + 2 * Et * cfg.num_classes
        own_bytes_step = sum(k["algo_bytes"] * k["launches"] for k in kernels.values()) / (min(K, 5) if args.cuda_graph else K)
        mem_bytes_step = own_bytes_step + batch * Cin * cfg.img_size ** 2 * 4 + nparam * 28
        tf_peak = peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1395.1))
        if cfg.precision != "bf16":
            tf_peak = tf_peak / 2      # fp32 configs run their GEMMs as TF32: half the bf16 tensor rate (nominal)
        t_gemm = gemm_flops_img * batch / (tf_peak * 1e12)
        t_mem = mem_bytes_step / (peaks["hbm_gbs"] * 1e9)
        step_s = ms_total / K / 1e3
        model_roof = {
            "gemm_gflop_per_image": round(gemm_flops_img / 1e9, 3), "memory_bound_MB_per_image": round(mem_bytes_step / batch / 1e6, 3),
            "tensor_peak_tflops": round(tf_peak, 1), "hbm_peak_gbs": peaks["hbm_gbs"],
            "gemm_only_roof_img_s": round(batch / t_gemm, 1), "serial_roof_img_s": round(batch / (t_gemm + t_mem), 1),
            "frac_of_gemm_roof": round(t_gemm / step_s, 4), "frac_of_serial_roof": round((t_gemm + t_mem) / step_s, 4),
            "note": "per GPU; serial roof = GEMMs at the sustained tensor peak + every memory-bound op at the HBM copy peak, no overlap",
        }
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu, _ = cpu_reference_leg(cfg, 3, 1, cpu_batch, args.config)
        out = {
            "metric": METRIC, "value": round(batch * world * K / (ms_total / 1e3), 1), "unit": "images/s",
            "n_gpus": world, "steps": K, "warmup": warm, "ms_per_step": round(ms_total / K, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if cfg.precision == "bf16" else "fp32", "data": "synthetic", "config": config,
            "e2e": {"value": round(batch * world * K / (ms_e2e / 1e3), 1), "unit": "images/s",
                    "h2d_bytes_per_step": (himg.numel() * himg.element_size() + hlab.numel() * hlab.element_size()) * world,
                    "d2h_bytes_per_step": 4 * world, "ms_per_step": round(ms_e2e / K, 3)},
            "gpu_launches": launches, "roofline": roof, "model_roofline": model_roof, "kernels": kernels,
            "cpu_baseline": cpu, "clocks": clocks,
            "final_loss": round(final_loss, 4),
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()
        if args.cuda_graph:
            # a live CUDA graph that captured NCCL kernels keeps the communicator busy: tearing the process group down
            # hung (observed: torchrun killed by its timeout after the JSON line).  Every rank is past the last
            # collective here, so leave without the teardown.
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
