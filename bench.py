#!/usr/bin/env python
"""bench.py - cGAN G+D train-step throughput (patches/s) on B200, BASELINE.json configs[1] per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--img-size S] [--batch B]

One "step" = the reference loop body (cgan/cgan_train_enhanced.py:297-369): n_critic=2 critic updates (two D
forwards + gradient penalty with double backward + clip + Adam each) and one generator update, on B synthetic paired
256x256 patches per GPU.  N>1: one rank per GPU (torchrun), batch sharded, gradient all-reduce over NCCL.

The JSON line (rank 0) carries
  value        patches/s with the step's inputs already resident in HBM: CUDA-graph replays of CGANTrainer.step
               (CUDA events, barrier + sync both sides, max over ranks); --no-graph steps eagerly instead
  e2e          the same step through the public API from PINNED HOST buffers: per-step H2D of the patches/boxes
               (train_step.HostPrefetcher: the copy of batch i+1 overlaps step i) and D2H of the 9 logged scalars,
               all inside the timed region
  roofline     the dominant tensor-core kernel family: algorithmic FLOPs per launch / CUDA-event duration per launch,
               measured live in a third, eager pass of the same K steps (events cannot bracket kernels inside a graph
               replay), against MEASURED_PEAKS.json (sustained bf16 TFLOP/s); traffic = ncu DRAM bytes per launch
  cpu_baseline the CPU oracle (oracle/cgan_oracle.py, a port of the reference's torch-CPU fp32 path) on a bounded
               sample of the workload, all host cores, rank 0 at N=1 only
`--impl reference` times that CPU oracle alone (the reference is Python/PyTorch and /root/reference does not travel
to the GPU box, so the restated port is the reference arm; oracle header explains how it is pinned).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "gan-calibrated-semi-supervised-learning_b200"))

METRIC = "cgan_train_patches_per_s"
UNIT = "patches/s"
N_CRITIC, LAMBDA_GP = 2, 1.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--img-size", type=int, default=256)
    ap.add_argument("--batch", type=int, default=64, help="patches per GPU")
    ap.add_argument("--cpu-sample", type=int, default=4, help="patches per CPU-baseline step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="step eagerly instead of replaying the captured CUDA graph")
    ap.add_argument("--mode", default="train", choices=["train", "infer"],
                    help="infer = BASELINE configs[4]: generator-only eval forward (cgan/inference.py:170-171), default batch 256")
    ap.add_argument("--refine", action="store_true",
                    help="the reference's real data flow: uint8 images resident in HBM, pred / gt patches letterboxed on the "
                         "device, the fake patch re-cropped from the generator's CURRENT delta in every critic step and in the "
                         "generator step (b200cgan.refine; cgan_train_enhanced.py:311-315,358-360) - the critic-step generator "
                         "forwards are live work in this mode")
    ap.add_argument("--config", default=None, choices=["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"],
                    help="BASELINE.json configs by number: cfg1 = B 4 S 256 (the CPU plumbing case), cfg2/cfg3 = B 64/GPU S 256 "
                         "(default; cfg3 is cfg2 under torchrun), cfg4 = B 16/GPU S 512, cfg5 = generator-only forward, B 256")
    a = ap.parse_args()
    if a.config == "cfg1":
        a.batch, a.img_size = 4, 256
    elif a.config == "cfg4":
        a.batch, a.img_size = 16, 512
    elif a.config == "cfg5":
        a.mode = "infer"
    return a


def workload_config(args, world):
    """The `config` block, identical for the B200 arm and the reference arm (the reference arm steps a bounded SAMPLE of this
    workload, described in its cpu_baseline.sample)."""
    S, B = args.img_size, args.batch
    name = {(64, 256): "configs[1]" if world == 1 else "configs[2]", (16, 512): "configs[3]", (4, 256): "configs[0]"}.get((B, S), "custom")
    return {"workload": f"BASELINE {name}: cGAN G+D train step (n_critic={N_CRITIC}, lambda_gp={LAMBDA_GP}, spectral norm, "
                        f"clip+Adam), bf16 storage / fp32 accumulate, batch {B} synthetic {S}x{S} paired patches per GPU",
            "img_size": S, "batch_per_gpu": B, "global_batch": B * world, "parallelism": f"dp{world}",
            "l2": "per-step working set (activation stash > 4 GB at B=64,S=256) far exceeds the 126 MB L2; two input "
                  "batches alternate"}


# --------------------------------------------------------------------------------------------- FLOP model
def conv_flops_per_patch(S):
    """(F_G, F_G1, F_D, F_D1): forward conv FLOPs per patch (MAC x 2), App. A of SURVEY.md."""
    def conv(cin, cout, hout):
        return 2.0 * hout * hout * cout * 16 * cin
    fg1 = conv(3, 64, S // 2)
    fg = fg1 + conv(64, 128, S // 4) + conv(128, 256, S // 8) + conv(256, 512, S // 16)
    for cin, cout, hin in ((512, 256, S // 16), (512, 128, S // 8), (256, 64, S // 4), (128, 64, S // 2)):
        fg += 2.0 * hin * hin * 16 * cin * cout
    fd1 = conv(6, 64, S // 2)
    fd = fd1 + conv(64, 128, S // 4) + conv(128, 256, S // 8) + conv(256, 512, S // 16)
    fd += 2.0 * (S // 16 - 1) ** 2 * 512 * 16
    return fg, fg1, fd, fd1


def useful_flops_per_patch(S, c=N_CRITIC, live_critic_g=False):
    """SURVEY.md §8d model.  With `refined` given as an input (every BASELINE config) the critic steps' no-grad generator
    forward has no consumer and is not executed, so its c * F_G is NOT counted; with a refine_fn it is live."""
    fg, fg1, fd, fd1 = conv_flops_per_patch(S)
    return ((c if live_critic_g else 0) + 3) * fg - fg1 + (12 * c + 1) * fd - 3 * c * fd1


GEMM_KERNELS = ("b2c_conv4s2_bf16", "b2c_convt4s2_bf16", "b2c_wgrad4s2_bf16")
# entry points that are variants of a family above (same kernels, same trailing size arguments): timed and reported with it
ALIASES = {"b2c_convt4s2_signbits_bf16": "b2c_convt4s2_bf16", "b2c_first_fprop_bits": "b2c_first_fprop"}
BW_KERNELS = ("b2c_in_act_fwd", "b2c_in_act_bwd", "b2c_in_act_bwd2", "b2c_in_act_pool_fwd", "b2c_first_fprop",
              "b2c_first_wgrad", "b2c_first_dgrad", "b2c_multi_adam", "b2c_lrelu_bwd", "b2c_head_fprop", "b2c_head_dgrad",
              "b2c_head_wgrad")


def bw_bytes(name, a):
    """Algorithmic bytes of one launch of a bandwidth kernel (each distinct tensor once) from its C-ABI
    argument pattern (include/b2c.h; scalars as given, pointers as True / False = given / NULL), or None."""
    if name == "b2c_in_act_fwd":         # z, ldz, zs, mask, slope, o0, ldo0, o1, ldo1, N, HW, C
        N, HW, C = a[-3:]
        return N * HW * C * (2 * (2 + (1 if a[7] else 0)) + (1 if a[3] else 0))
    if name == "b2c_in_act_pool_fwd":
        N, HW, C = a[-3:]
        return N * HW * C * 2
    if name == "b2c_in_act_bwd":         # g, ldg, gvec, fwd_sums, z, ldz, zs, mask, slope, ws, gz, ldgz, N, HW, C
        N, HW, C = a[-3:]
        return N * HW * C * (2 * (3 if a[0] else 2) + (1 if a[7] else 0))
    if name == "b2c_in_act_bwd2":
        N, HW, C = a[-3:]
        return N * HW * C * 2 * 5
    if name == "b2c_lrelu_bwd":          # g, a, out, n, slope
        return a[3] * 2 * 3
    if name == "b2c_first_fprop":        # a0, a1, b0, b1, alpha, w, inv_sigma, bias, y, N, H, W, slope[, sign_bits]
        N, H, W = a[9:12]
        bits = 8 if len(a) > 13 else 0   # b2c_first_fprop_bits: one 64-bit word per output pixel
        return N * 3 * H * W * 4 * (2 if a[1] else 1) * (2 if a[2] else 1) + N * (H // 2) * (W // 2) * (64 * 2 + bits)
    if name == "b2c_first_wgrad":        # a0, a1, b0, b1, alpha, gz, dw, db, scale, wdot, part, nparts, N, H, W
        N, H, W = a[-3:]
        return N * 3 * H * W * 4 * (2 if a[1] else 1) * (2 if a[2] else 1) + N * (H // 2) * (W // 2) * 64 * 2
    if name == "b2c_first_dgrad":        # gz, w, inv_sigma, gx0, gx1, ws, N, H, W
        N, H, W = a[-3:]
        return N * 3 * H * W * 4 * (2 if a[4] else 1) + N * (H // 2) * (W // 2) * 64 * 2
    if name in ("b2c_head_fprop", "b2c_head_dgrad", "b2c_head_wgrad"):
        N, h, C = a[-3:]
        return N * h * h * C * 2
    return None


def gemm_flops(name, a):
    """Algorithmic FLOPs of one launch from its C-ABI argument pattern (include/b2c.h; trailing arguments are the sizes)."""
    if name == "b2c_wgrad4s2_bf16":      # ..., N, Hs, Ws, Cs, Cl, sm_count
        N, Hs, Ws, Cs, Cl = a[-6:-1]
        return 2.0 * N * Hs * Ws * Cs * 16 * Cl
    N, H, W, Cin, Cout = a[-5:]           # ..., N, H, W, Cin, Cout
    if name == "b2c_conv4s2_bf16":
        return 2.0 * N * (H // 2) * (W // 2) * Cout * 16 * Cin
    return 2.0 * N * H * W * Cout * 16 * Cin


def kernel_families(timed, K, adam_bytes):
    """Aggregate the CUDA-event records of _lib.time_kernels over K steps -> (tensor-core families, bandwidth families)."""
    timed = dict(timed)
    for alias, base in ALIASES.items():
        timed[base] = list(timed.get(base, [])) + list(timed.pop(alias, []))
    bwfam = {}
    for name in BW_KERNELS:
        recs = timed.pop(name, [])
        if not recs:
            continue
        ms = sum(a.elapsed_time(b) for a, b, _ in recs)
        if name == "b2c_multi_adam":  # launch order within a step: D, D, G
            nb = sum(adam_bytes[i % len(adam_bytes)] for i in range(len(recs)))
        else:
            nb = sum(bw_bytes(name, sc) or 0 for _, _, sc in recs)
        bwfam[name] = {"launches_per_step": len(recs) / K, "ms_per_step": ms / K, "gbs": nb / ms / 1e6 if ms else 0.0,
                       "bytes_per_launch": nb / len(recs), "ms_per_launch": ms / len(recs)}
    fam = {}
    for name, recs in timed.items():
        if not recs:
            continue
        ms = sum(a.elapsed_time(b) for a, b, _ in recs)
        fl = sum(gemm_flops(name, sc) for _, _, sc in recs)
        fam[name] = {"launches_per_step": len(recs) / K, "ms_per_step": ms / K, "tflops": fl / ms / 1e9 if ms else 0.0,
                     "flop_per_launch": fl / max(len(recs), 1), "ms_per_launch": ms / max(len(recs), 1)}
    return fam, bwfam


def roofline_blocks(fam, bwfam, ms_step):
    """-> (`roofline`: dominant tensor-core family vs the measured sustained bf16 peak, `roofline_bw`: dominant bandwidth
    kernel vs the measured HBM copy bandwidth), both from durations measured live with CUDA events."""
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    peak_tf, peak_src = peaks.get("bf16_tflops_sustained", 1400.0), ("measured" if peaks else "fallback")
    peak_bw = peaks.get("hbm_gbs", 6500.0)
    roofline = None
    if fam:
        dom = max(fam, key=lambda n: fam[n]["ms_per_step"])
        total_ms = sum(f["ms_per_step"] for f in fam.values())
        total_tf = sum(f["tflops"] * f["ms_per_step"] for f in fam.values()) / total_ms if total_ms else 0.0
        traffic = None  # ncu dram__bytes_read+write per launch of the dominant family (one --set full capture of a step)
        for tfn in ("r02f_traffic.json", "r02_traffic.json", "r01_traffic.json"):
            tf = ROOT / "profiles" / tfn
            if tf.exists():
                traffic = json.loads(tf.read_text()).get(dom, {}).get("dram_bytes_per_launch")
                break
        roofline = {"bound": "tensor", "kernel": dom, "achieved": fam[dom]["tflops"], "peak": peak_tf, "unit": "TFLOP/s",
                    "frac": fam[dom]["tflops"] / peak_tf, "traffic": traffic, "peak_source": peak_src + " bf16_tflops_sustained",
                    "flop_per_launch": fam[dom]["flop_per_launch"], "ms_per_launch": fam[dom]["ms_per_launch"],
                    "share_of_step": fam[dom]["ms_per_step"] / ms_step,
                    "all_tensor_kernels": {"tflops": total_tf, "frac": total_tf / peak_tf, "share_of_step": total_ms / ms_step}}
    roofline_bw = None
    if bwfam:
        bdom = max(bwfam, key=lambda n: bwfam[n]["ms_per_step"])
        bms = sum(f["ms_per_step"] for f in bwfam.values())
        bgbs = sum(f["gbs"] * f["ms_per_step"] for f in bwfam.values()) / bms if bms else 0.0
        roofline_bw = {"bound": "hbm", "kernel": bdom, "achieved": bwfam[bdom]["gbs"], "peak": peak_bw, "unit": "GB/s",
                       "frac": bwfam[bdom]["gbs"] / peak_bw, "peak_source": peak_src + " hbm_gbs",
                       "bytes_per_launch": bwfam[bdom]["bytes_per_launch"], "ms_per_launch": bwfam[bdom]["ms_per_launch"],
                       "share_of_step": bwfam[bdom]["ms_per_step"] / ms_step,
                       "all_bandwidth_kernels": {"gbs": bgbs, "frac": bgbs / peak_bw, "share_of_step": bms / ms_step},
                       "per_kernel": {n: {"gbs": round(f["gbs"], 1), "frac": round(f["gbs"] / peak_bw, 3),
                                          "ms_per_step": round(f["ms_per_step"], 4),
                                          "launches_per_step": f["launches_per_step"]} for n, f in bwfam.items()}}
    return roofline, roofline_bw


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v == "Active":
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------- CPU oracle legs
def cpu_oracle_run(S, sample, steps, warmup):
    """Time `steps` train steps of the CPU oracle on `sample` patches, all host cores.  -> (patches/s, s/step, cores)."""
    import torch
    sys.path.insert(0, str(ROOT / "oracle"))
    import cgan_oracle as O  # the checker / CPU baseline; never on the product path
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    st = O.init_state(42)
    times = []
    for i in range(warmup + steps):
        batch, masks, alphas = O.make_batch(sample, S, i)
        t0 = time.perf_counter()
        O.train_step(st, batch, masks, alphas, n_critic=N_CRITIC, lambda_gp=LAMBDA_GP)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    return sample * steps / total, total / steps, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    S, sample = args.img_size, args.cpu_sample
    v, spstep, cores = cpu_oracle_run(S, sample, args.steps, args.warmup)
    desc = (f"{sample} synthetic {S}x{S} paired patches per step (of {args.batch}/GPU in the GPU workload), fp32, "
            f"torch CPU with {cores} threads, {args.steps} steps after {args.warmup} warm-up")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": spstep * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, max(args.gpus, 1)),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- the B200 arm
def synth_host_batches(B, S, rank, count=2):
    """Pinned host batches (SURVEY.md §8d distributions); `count` distinct batches are rotated over the steps."""
    import torch
    out = []
    for i in range(count):
        g = torch.Generator("cpu").manual_seed(1234 + 1000 * rank + i)
        u = lambda *s: torch.rand(*s, generator=g)
        b = {"pred": u(B, 3, S, S) * 2 - 1, "gt": u(B, 3, S, S) * 2 - 1, "refined": u(B, 3, S, S) * 2 - 1,
             "pred_box": torch.cat([u(B, 2) * 0.6 + 0.2, u(B, 2) * 0.3 + 0.1], 1),
             "delta_true": (torch.randn(B, 4, generator=g) * torch.tensor([0.17, 0.23, 0.32, 0.32])).clamp(-1.5, 1.5)}
        out.append({k: v.pin_memory() for k, v in b.items()})
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from b200cgan import _lib, rng
    from b200cgan.models import Discriminator, GeneratorUNet, weights_init_normal
    from b200cgan.train_step import CGANTrainer

    S, B, K, W = args.img_size, args.batch, args.steps, args.warmup
    torch.manual_seed(42)
    G, D = GeneratorUNet(0.3), Discriminator(True)
    G.apply(weights_init_normal); D.apply(weights_init_normal)
    G, D = G.to(dev), D.to(dev)
    rng.manual_seed(42)  # one stream for the whole job; each rank reads the slice of its global sample indices (rng.set_shard)
    host = synth_host_batches(B, S, rank)
    resident = [{k: v.to(dev) for k, v in b.items()} for b in host]
    refine_fn = None
    if args.refine:
        from b200cgan import refine
        gi = torch.Generator("cpu").manual_seed(77 + rank)
        bank = refine.DeviceImageBank([torch.randint(0, 256, (480, 640, 3), dtype=torch.uint8, generator=gi) for _ in range(32)], dev)
        sample_img = torch.randint(0, 32, (B,), generator=gi).to(dev, torch.int32)
        box_buf = resident[0]["pred_box"].clone()  # persistent: the captured graph's refine kernel reads it
        for b in resident:  # patches come from the images; both batches share the boxes the refine kernel sees
            b["pred_box"] = box_buf
            b["pred"] = refine.letterbox_patches(bank, sample_img, box_buf, S)
            b["gt"] = refine.letterbox_patches(bank, sample_img, (box_buf * 1.02).clamp(0.05, 0.9), S)
            b["refined"] = None
        for hb, b in zip(host, resident):
            hb["pred"], hb["gt"] = b["pred"].cpu().pin_memory(), b["gt"].cpu().pin_memory()
            hb["pred_box"] = box_buf.cpu().pin_memory()
            hb["refined"] = None
        refine_fn = refine.make_refine_fn(bank, sample_img, box_buf, S, fallback=None)
    tr = CGANTrainer(G, D, n_critic=N_CRITIC, lambda_gp=LAMBDA_GP, world_size=world, rank=rank if world > 1 else None,
                     use_cuda_graph=not args.no_graph, refine_fn=refine_fn)
    keys = ("pred", "gt", "refined", "pred_box", "delta_true")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    # ---- device-resident leg: `value`.  With CUDA-graph mode the first two steps run eagerly, the third is captured.
    tr.step(*[resident[0][k] for k in keys])  # first step: allocates the weight-pack buffers (lazy per-weight packs)
    l0 = _lib.launch_count()
    tr.step(*[resident[1][k] for k in keys])
    launches_per_step = _lib.launch_count() - l0  # libb2c entry-point calls of one steady-state (eager) step
    for i in range(2, max(W, 3)):
        tr.step(*[resident[i % 2][k] for k in keys])
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        out = tr.step(*[resident[i % 2][k] for k in keys])
    e1.record()
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1))
    launches = launches_per_step * K
    clk = clocks.stop() if rank == 0 else None
    finite = all(bool(torch.isfinite(v).all()) for v in out.scalars.values())

    # ---- end-to-end leg: pinned host -> device every step, scalars back to the host every step
    h2d = sum(v.numel() * v.element_size() for v in host[0].values() if v is not None)
    from b200cgan.train_step import HostPrefetcher
    live = [k for k in keys if host[0][k] is not None]
    pf = HostPrefetcher([host[0][k] for k in live], dev)  # copy of batch i+1 overlaps step i (side stream)

    def e2e_steps(n):
        nbytes = 0
        pf.submit([host[0][k] for k in live])
        for i in range(n):
            if i + 1 < n:
                pf.submit([host[(i + 1) % 2][k] for k in live])
            got = dict(zip(live, pf.take()))
            o = tr.step(*[got.get(k) for k in keys])
            pf.release()
            nbytes = 4 * len(o.to_host())  # the 9 logged scalars come back to the host every step
        return nbytes

    e2e_steps(2)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    d2h = e2e_steps(K)
    f1.record()
    barrier()
    ms_e2e = max_over_ranks(f0.elapsed_time(f1))

    # ---- roofline leg: the same K steps run eagerly with CUDA events around every tensor-core entry point
    tr.disable_cuda_graph()
    tr.step(*[resident[0][k] for k in keys])
    barrier()
    timed = _lib.time_kernels(GEMM_KERNELS + BW_KERNELS + tuple(ALIASES))
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for i in range(K):
        tr.step(*[resident[i % 2][k] for k in keys])
    g1.record()
    barrier()
    ms_eager = g0.elapsed_time(g1)
    _lib.time_kernels(None)
    fam, bwfam = kernel_families(timed, K, [28 * tr.optD.total] * N_CRITIC + [28 * tr.optG.total])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    roofline, roofline_bw = roofline_blocks(fam, bwfam, ms_dev / K)
    value = world * B * K / (ms_dev / 1e3)
    useful = useful_flops_per_patch(S, live_critic_g=args.refine)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": world * B * K / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / K},
            "gpu_launches": launches, "cuda_graph": not args.no_graph, "eager_ms_per_step": ms_eager / K,
            "refine": ("device: fake patches re-cropped from G's current delta (3 crops + 2 extra generator forwards per step)"
                       if args.refine else "input tensor (BASELINE configs: synthetic paired patches)"),
            "roofline": roofline, "roofline_bw": roofline_bw, "kernels": fam, "clocks": clk,
            "useful_conv_tflops": useful * B / (ms_dev / K) / 1e9, "finite": finite}
    if world == 1 and not args.no_cpu_baseline:
        v, spstep, cores = cpu_oracle_run(S, args.cpu_sample, 2, 1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{args.cpu_sample} patches {S}x{S} per step, 2 steps after 1 warm-up, fp32 "
                                          f"torch CPU oracle ({spstep:.2f} s/step)"}
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_infer(args):
    """BASELINE configs[4]: generator-only eval forward on B synthetic patches (pseudo-label calibration throughput).
    One line, same keys; `value` = patches/s with inputs resident, `e2e` = through b200cgan.inference.calibrate_patches
    (pinned host patches in, (B,4) deltas back on the host)."""
    import torch
    if args.impl == "reference":
        sys.path.insert(0, str(ROOT / "oracle"))
        import cgan_oracle as O
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        st = O.init_state(42)
        x = torch.rand(args.cpu_sample, 3, args.img_size, args.img_size) * 2 - 1
        ts = []
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            with torch.no_grad():
                O.g_forward(st["G"], x, 0.3, train=False)
            if i >= args.warmup:
                ts.append(time.perf_counter() - t0)
        v = args.cpu_sample * len(ts) / sum(ts)
        print(json.dumps({"impl": "reference", "metric": "cgan_infer_patches_per_s", "value": v, "unit": UNIT,
                          "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
                          "dtype": "f32", "data": "synthetic",
                          "config": {"workload": f"G eval forward, {args.img_size}x{args.img_size}, CPU sample of {args.cpu_sample}"},
                          "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                           "sample": f"{args.cpu_sample} patches per forward"},
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)
        return
    from b200cgan import _lib, inference
    from b200cgan.models import GeneratorUNet, weights_init_normal
    S, B, K, W = args.img_size, (args.batch if args.batch != 64 else 256), args.steps, max(args.warmup, 3)
    torch.manual_seed(42)
    G = GeneratorUNet(0.3)
    G.apply(weights_init_normal)
    G = G.cuda().eval()
    host = [(torch.rand(B, 3, S, S, generator=torch.Generator().manual_seed(i)) * 2 - 1).pin_memory() for i in range(2)]
    dev = [h.cuda() for h in host]
    with torch.no_grad():
        for i in range(W):
            G(dev[i % 2])
        torch.cuda.synchronize()
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(K):
            d = G(dev[i % 2])
        e1.record()
        torch.cuda.synchronize()
        launches = _lib.launch_count() - l0
        ms = e0.elapsed_time(e1)
        inference.calibrate_patches(G, host[0], B)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(K):
            out = inference.calibrate_patches(G, host[i % 2], B)
        f1.record()
        torch.cuda.synchronize()
        ms2 = f0.elapsed_time(f1)
        timed = _lib.time_kernels(GEMM_KERNELS + BW_KERNELS + tuple(ALIASES))
        for i in range(K):
            G(dev[i % 2])
        torch.cuda.synchronize()
        _lib.time_kernels(None)
    fam, bwfam = kernel_families(timed, K, [0])
    roofline, roofline_bw = roofline_blocks(fam, bwfam, ms / K)
    cpu = None
    if not args.no_cpu_baseline:
        sys.path.insert(0, str(ROOT / "oracle"))
        import cgan_oracle as O
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        st = O.init_state(42)
        xs = torch.rand(args.cpu_sample * 4, 3, S, S) * 2 - 1
        with torch.no_grad():
            O.g_forward(st["G"], xs, 0.3, train=False)
            t0 = time.perf_counter()
            for _ in range(2):
                O.g_forward(st["G"], xs, 0.3, train=False)
            dt = (time.perf_counter() - t0) / 2
        cpu = {"value": xs.shape[0] / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{xs.shape[0]} patches {S}x{S} per forward, 2 forwards after 1 warm-up, fp32 torch CPU oracle"}
    fg = conv_flops_per_patch(S)[0]
    print(json.dumps({"metric": "cgan_infer_patches_per_s", "value": B * K / (ms / 1e3), "unit": UNIT, "n_gpus": 1,
                      "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                      "config": {"workload": f"BASELINE configs[4]: GeneratorUNet eval forward, batch {B} synthetic {S}x{S} patches",
                                 "img_size": S, "batch_per_gpu": B, "l2": "two input batches alternate; activations (>1 GB) exceed L2"},
                      "e2e": {"value": B * K / (ms2 / 1e3), "unit": UNIT, "h2d_bytes_per_step": host[0].numel() * 4,
                              "d2h_bytes_per_step": B * 16, "ms_per_step": ms2 / K},
                      "gpu_launches": launches, "useful_conv_tflops": fg * B / (ms / K) / 1e9,
                      "roofline": roofline, "roofline_bw": roofline_bw, "kernels": fam, "cpu_baseline": cpu,
                      "finite": bool(torch.isfinite(d).all()) and bool(torch.isfinite(out).all())}), flush=True)


if __name__ == "__main__":
    a = parse()
    # The contract is ONE JSON line on stdout.  Libraries (NCCL prints its version banner to stdout) must not share it:
    # file descriptor 1 is pointed at stderr for the run and the JSON line goes to the saved original.
    sys.stdout.flush()
    _json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _json_out
    if a.mode == "infer":
        if int(os.environ.get("RANK", "0")) == 0:
            run_infer(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
