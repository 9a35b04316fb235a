#!/usr/bin/env python
"""bench.py -- headline benchmark of the DSP front end (BASELINE.json: audio-sec/sec, features + mel + aug).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload features_mel|features_mel_aug]
  python bench.py --impl reference ...      # the reference's CPU path (oracle port) on the host cores

A "step" is one pass of the hot path over one batch of synthetic input: 72 clips (24 triplets) of
10 s / 44.1 kHz / 8 channels per GPU (BASELINE.json configs[1]; configs[2] with --workload
features_mel_aug adds the augmentation chain that makes the 24 negatives).  Under torchrun every
rank processes its own 72 clips (weak scaling, no data-path collective); the all-gather into the
global contrastive batch is timed separately and reported under "allgather".
Prints ONE JSON line on rank 0.
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

SR, CLIP_S, T = 44100, 10.0, 441000
N_CLIPS = 72                           # 24 triplets: anchor / positive / negative
N_MELS, HOP, NFFT = 128, 256, 1024
F = 1 + T // HOP
ALG_BYTES_PER_CLIP = 8 * T * 4 + 8 * N_MELS * F * 4 + 64 * 4          # 21 169 664 B
METRIC = "audio-sec/sec features+mel+aug"
UNIT = "audio-s/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default=os.environ.get("MIXFE_BENCH_WORKLOAD", "features_mel"),
                    choices=["features_mel", "features_mel_aug"])
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--ref-clips", type=int, default=0, help="clips per reference step (0 = one per worker, max 24)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def config_dict(workload, n_gpus, extra=None):
    c = {"workload": ("24-triplet batch (72 clips x 10 s, 44.1 kHz, 8 ch) features + 8ch 128-mel log-mel"
                      + (" + augmentation chain for the 24 negatives" if workload == "features_mel_aug" else "")),
         "clips_per_gpu": N_CLIPS, "clip_seconds": CLIP_S, "sample_rate": SR, "n_fft": NFFT, "hop": HOP,
         "n_mels": N_MELS, "parallelism": f"clip-sharded dp{n_gpus}",
         "l2": "inputs (1.02 GB per step) exceed the 126 MB L2; no explicit flush"}
    if extra:
        c.update(extra)
    return c


# =================================================================================================
# reference arm: the reference's own CPU algorithm (oracle port; the reference is pure Python on
# torch/torchaudio/scipy, restated in oracle/ and pinned bit-exactly against it)
# =================================================================================================
_WORKER_CLIPS = None


def _ref_worker_init(seed_base, n_local, workload):
    global _WORKER_CLIPS
    import torch
    torch.set_num_threads(1)
    from oracle import mixfe_oracle as O
    _WORKER_CLIPS = [O.synthetic_clip(seed_base + i, T) for i in range(n_local)]
    O.MelFrontEnd(O.FrontEndConfig())          # warm the filterbank code path


def _ref_worker_run(job):
    idx, workload = job
    import torch
    from oracle import mixfe_oracle as O
    clip = _WORKER_CLIPS[idx % len(_WORKER_CLIPS)]
    t0 = time.perf_counter()
    if workload == "features_mel_aug" and idx % 3 == 2:          # every third clip is a negative
        torch.manual_seed(1234 + idx)
        clip = O.apply_clip_aug(clip, O.draw_clip_aug(prob=0.5), fft_reverb=True)
    f = O.mixing_features(clip)
    lm = O.log_mel({n: v[None] for n, v in clip.items()})
    return time.perf_counter() - t0, float(f.sum()) + float(lm[0, 0, 1, 0])


_REAL_STDOUT = None


def emit(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    import torch
    torch.set_num_threads(1)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    workers = max(1, min(cores, 64))
    clips = args.ref_clips or min(N_CLIPS, 2 * workers)
    workers = min(workers, clips)
    ctx = mp.get_context("fork")
    pool = ctx.Pool(workers, initializer=_ref_worker_init, initargs=(100, 2, args.workload))
    jobs = [(i, args.workload) for i in range(clips)]
    try:
        for _ in range(max(1, args.warmup)):
            pool.map(_ref_worker_run, jobs, chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_ref_worker_run, jobs, chunksize=1)
        dt = time.perf_counter() - t0
    finally:
        pool.terminate()
    ms = dt / args.steps * 1e3
    value = clips * CLIP_S / (ms / 1e3)
    sample = (f"{clips} clips x 10 s per step ({clips * CLIP_S:.0f} audio-s), {workers} worker processes x 1 torch thread "
              f"(the reference's DataLoader-worker deployment, scripts/train_baseline.sh:13-19)")
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": max(1, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": config_dict(args.workload, args.gpus, {"reference_sample_clips_per_step": clips}),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample,
                             "host_cores_visible": cores},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# =================================================================================================
# native arm
# =================================================================================================
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def mark_begin(self):
        self.t_begin = time.perf_counter()

    def mark_end(self):
        self.t_end = time.perf_counter()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t_b, t_e = getattr(self, "t_begin", 0.0), getattr(self, "t_end", float("inf"))
        inside = [ln for (ts, ln) in self.lines if t_b <= ts <= t_e + 0.05]
        if not inside:                       # timed region shorter than one sampling period: nearest samples
            inside = [ln for (ts, ln) in self.lines if ts >= t_b - 0.2][:3] or [ln for (_, ln) in self.lines[-3:]]
        for ln in inside:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); smax.append(float(p[2])); power.append(float(p[3]))
            except ValueError:
                continue
            for nme, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power)}


def measured_peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_baseline_subprocess(workload):
    """The oracle timed on this box's host cores, before CUDA is initialised in this process."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2", "--warmup", "1",
           "--workload", workload]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
        for ln in reversed(out.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": "failed: " + out.stderr[-300:]}
    except Exception as e:                                                  # noqa: BLE001
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e}"}


def run_native(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline_subprocess(args.workload)

    import torch
    import torch.distributed as dist
    import mixing_style_transfer_b200 as pkg
    from mixing_style_transfer_b200.distributed import GlobalBatch
    from mixing_style_transfer_b200.pipeline import HostPipeline

    if not torch.cuda.is_available():
        raise SystemExit("bench.py (native arm) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    # ---- synthetic batch, HBM-resident --------------------------
    g = torch.Generator(device=dev).manual_seed(20261104 + rank)
    stems = []
    for _ in pkg.STEMS:
        gain = 10 ** (-12.0 * torch.rand(1, device=dev, generator=g) / 20)
        stems.append((torch.rand(N_CLIPS, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 * gain)
    fe = pkg.MixingFeatureExtractor()
    plan = fe._plan(dev)
    aug = pkg.AudioAugmenter(prob=0.5)
    with_aug = args.workload == "features_mel_aug"
    decisions = None
    if with_aug:
        torch.manual_seed(1234)
        decisions = [aug.draw(impulse_device=dev) for _ in range(N_CLIPS // 3)]
        aug_plan = aug._plan(dev)
    gb = GlobalBatch(N_CLIPS, plan.feature_dim, N_MELS, F, dev, with_logmel=True)
    f_out, lm_out = gb.local_views()

    def step():
        if with_aug:
            # negatives (clips 48..71) = augment_stems(anchors 0..23), written in place into the batch
            anchors = {n: s[0:N_CLIPS // 3] for n, s in zip(pkg.STEMS, stems)}
            aug.apply(anchors, decisions, out={n: s[2 * N_CLIPS // 3:] for n, s in zip(pkg.STEMS, stems)})
        plan.features_logmel(stems, True, True, features_out=f_out, logmel_out=lm_out)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    plan.enable_timing(True)
    if with_aug:
        aug_plan.enable_timing(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    sampler.mark_begin()
    n0 = pkg.launch_count()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    sampler.mark_end()
    barrier()
    launches = pkg.launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    ms = e0.elapsed_time(e1) / args.steps
    k1_ms, k1_n = plan.kernel_time("features")
    k2_ms, k2_n = plan.kernel_time("finalize")
    plan.enable_timing(False)
    kernels = {"k1_features_ms": k1_ms / max(1, k1_n), "k2_finalize_ms": k2_ms / max(1, k2_n)}
    if with_aug:
        a_ms, a_n = aug_plan.kernel_time("augment")
        r_ms, r_n = aug_plan.kernel_time("reverb")
        aug_plan.enable_timing(False)
        kernels.update({"k3_augment_ms": a_ms / max(1, a_n), "k4_reverb_ms_per_step": r_ms / max(1, args.steps)})
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * N_CLIPS * CLIP_S / (ms / 1e3)

    # ---- all-gather into the global contrastive batch, timed on its own -----
    allgather = None
    if world > 1:
        ggb = GlobalBatch(N_CLIPS, plan.feature_dim, N_MELS, F, dev, with_logmel=True)
        gf, glm = ggb.local_views()
        gf.copy_(f_out); glm.copy_(lm_out)
        for _ in range(2):
            ggb.all_gather()
        torch.cuda.synchronize(); barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        reps = 5
        for _ in range(reps):
            ggb.all_gather()
        a1.record(); torch.cuda.synchronize(); barrier()
        t = torch.tensor([a0.elapsed_time(a1) / reps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ag_ms = float(t.item())
        allgather = {"ms": ag_ms, "bytes_received_per_rank": ggb.payload_bytes_received(),
                     "recv_GBs_per_rank": ggb.payload_bytes_received() / ag_ms / 1e6,
                     "note": "in-place all_gather_into_tensor of (72,64) features + (72,8,128,1723) fp32 log-mels per rank; "
                             "not part of value (extraction is collective-free)"}
        del ggb
        # config 5 (global contrastive batch): extraction followed by the gather (NCCL), the two pipelined in 4 chunks
        # (gather of chunk c on a side stream under the kernels of chunk c + 1), and the same with copy-engine pushes
        # into symmetric (NVLink-mapped) buffers, which need no SMs and therefore do overlap the persistent kernel
        from mixing_style_transfer_b200.distributed import extract_global_batch
        stems_d = dict(zip(pkg.STEMS, stems))
        comm = torch.cuda.Stream(device=dev)
        both = {}
        ref_gb = extract_global_batch(fe, stems_d, n_chunks=4, comm_stream=comm)
        torch.cuda.synchronize()
        for label, nch, sym in (("nccl_serial_ms", 1, False), ("nccl_pipelined_4_chunks_ms", 4, False),
                                ("symmetric_ce_serial_ms", 1, True), ("symmetric_ce_pipelined_4_chunks_ms", 4, True),
                                ("multicast_ce_serial_ms", 1, "multicast"), ("multicast_ce_pipelined_4_chunks_ms", 4, "multicast")):
            gbx = GlobalBatch(N_CLIPS, plan.feature_dim, N_MELS, F, dev, with_logmel=True, n_chunks=nch, symmetric=sym)
            for _ in range(2):
                extract_global_batch(fe, stems_d, n_chunks=nch, comm_stream=comm, out=gbx)
            torch.cuda.synchronize(); barrier()
            a0.record()
            for _ in range(reps):
                extract_global_batch(fe, stems_d, n_chunks=nch, comm_stream=comm, out=gbx)
            a1.record(); torch.cuda.synchronize(); barrier()
            t = torch.tensor([a0.elapsed_time(a1) / reps], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            both[label] = float(t.item())
            if nch == 4:
                both[label.replace("_ms", "_equals_nccl")] = bool(torch.equal(gbx.logmel, ref_gb.logmel)
                                                                  and torch.equal(gbx.features, ref_gb.features))
            del gbx
        del ref_gb
        allgather["extract_plus_gather"] = both

    # ---- end to end from pinned host memory through the public pipeline ---------------------------
    e2e = None
    if not args.no_e2e:
        host = {n: HostPipeline.pinned_like((N_CLIPS, 2, T)) for n in pkg.STEMS}
        for n, s in zip(pkg.STEMS, stems):
            host[n].copy_(s)
        hf = HostPipeline.pinned_like((N_CLIPS, plan.feature_dim))
        hl = HostPipeline.pinned_like((N_CLIPS, 8, N_MELS, F))
        pipe = HostPipeline(fe, N_CLIPS, T, chunk=12, device=dev, with_logmel=True)
        pipe.run(host, hf, hl)
        torch.cuda.synchronize(); barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            pipe.run(host, hf, hl)
        torch.cuda.synchronize()
        e_ms = (time.perf_counter() - t0) / args.e2e_steps * 1e3
        if world > 1:
            t = torch.tensor([e_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_ms = float(t.item())
        ok = bool(torch.allclose(hf, f_out.cpu(), rtol=0, atol=0)) if not with_aug else True
        e2e = {"value": world * N_CLIPS * CLIP_S / (e_ms / 1e3), "unit": UNIT, "ms_per_step": e_ms,
               "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes,
               "steps": args.e2e_steps, "matches_resident_run": ok,
               "api": "HostPipeline.run: pinned host stems -> chunked H2D -> features_and_logmel kernels -> D2H of "
                      "features and log-mels (features + mel only; augmentation decisions are host-side)"}

    if rank == 0:
        peak, peak_src = measured_peak_hbm()
        k1 = kernels["k1_features_ms"]
        achieved = N_CLIPS * ALG_BYTES_PER_CLIP / (k1 / 1e3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "k1_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_dict(args.workload, world),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "kernel": "k1_features", "kernel_ms": k1, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": N_CLIPS * ALG_BYTES_PER_CLIP,
                             "note": "K1 is FP32-issue-bound (20 FLOP/B), the HBM fraction is what BASELINE.json asks for"},
                "kernels": kernels, "gpu_launches": launches, "clocks": clocks}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if e2e is not None:
            line["e2e"] = e2e
        if allgather is not None:
            line["allgather"] = allgather
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner with
    # printf at communicator creation), so fd 1 is pointed at stderr for the run and the line goes to the real stdout.
    a = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
