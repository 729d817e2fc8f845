#!/usr/bin/env python
"""Headline benchmark: train samples/s (+ greedy captions/s) of the AttemptFour lc_NIC hot path on
synthetic NSD-shaped data (BASELINE.json config[0]: single subject, 327 684-vertex betas, LSTM 512,
vocab 5001, max len 15, batch 64 per GPU), next to the host-CPU restatement of the reference.

  python bench.py --gpus N --steps K --warmup W            # our arm (one rank per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle port, all host threads

Prints ONE JSON line (rank 0).  See DESIGN.md §measurement for what every key means.
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train samples/sec"
WORKLOAD = "C1 lc_NIC single subject: B=64/GPU, N=327684 betas, R=345 parcels x32, LSTM 512, vocab 5001, T=15, fp32"
WORKLOAD_C2 = ("C2 ThinkAndTell att_model variant of the C1 decoder (plain Bahdanau projections, pad-masked sparse CE "
               "from logits, teacher forced): B=64/GPU, N=327684 betas, R=345 parcels x32, LSTM 512, vocab 5001, T=15, fp32")
ALG_BYTES_PER_SAMPLE = 11.7e6      # 746 MB algorithmic HBM traffic per 64-sample step
WORKLOAD_C4 = ("C4 ms2_NIC two subjects (main_ms.py): two encoders x 360 parcels x32, shared decoder, 64 rows per subject = "
               "128 rows/GPU/step, N=327684 betas, LSTM 512, vocab 5001, T=15, fp32")
# C4 per 128-row step: X twice 2x168 MB, two packed encoder kernels 2x41 MB, other weights 49 MB, gradient write 102 MB,
# optimizer 28 B x 25.6 M = 717 MB -> 1.29 GB = 10.1 MB per sample
ALG_BYTES_PER_SAMPLE_C4 = 10.1e6


def workload_name(w):
    return {"C1": WORKLOAD, "C2": WORKLOAD_C2, "C4": WORKLOAD_C4}[w]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)      # 0.2 s timed region: resolves 1 % changes of the 1 ms step
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dropout", action="store_true")
    ap.add_argument("--workload", default="C1", choices=["C1", "C2", "C4"],
                    help="C1: BASELINE.json configs[0] (lc_NIC from config.yaml).  C2: configs[1], the ThinkAndTell "
                         "att_model variant of the same decoder (bare attention projections, pad-masked sparse CE "
                         "on logits): identical shapes and traffic.  C4: configs[3], the two-subject model of main_ms.py "
                         "(--batch = rows per subject)")
    return ap.parse_args()


FP32_FMA_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12     # 74.4: 128 FMA lanes per SM at the 1965 MHz boost clock
# dram__bytes_read.sum + dram__bytes_write.sum per launch from profiles/ (ncu --set full), by kernel group
NCU_TRAFFIC = {   # profiles/r1g_ncu_step.md (default streams, one steady-state step)
    "opt_adam_kernel (encoder part)": 206.0e6,    # 149.4 MB read + 56.6 MB written (algorithmic 261.4 MB: w, g partly in L2)
    "opt_adam_kernel (decoder part)": 118.0e6,    # 98.6 + 19.3 MB (algorithmic 172.5 MB)
    "opt_norm_kernel (decoder part)": 49.3e6,
    "opt_adam_kernel": 379.3e6,                   # whole vector, profiles/r1e_ncu_step.md
    "opt_norm_kernel": 127.6e6,
    "enc_transpose_kernel": 110.5e6,
    "enc_fwd_t": 117.4e6,
    "enc_bwd_dw_t": 88.7e6,
}


def tensor_peak():
    """Measured dense bf16 TFLOP/s (MEASURED_PEAKS.json, sustained figure: kernels are timed inside the step)."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d.get("bf16_tflops_sustained") or d.get("bf16_tflops"))
    except Exception:
        return 1590.0


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe).  A separate process
    (`nvidia-smi -lms 50`), started before the warm-up so that it is already streaming when the timed region begins;
    every line is stamped on arrival and only the samples inside [begin, end] are reported (all of them if the region
    was shorter than one period).  In-process NVML polling was tried and dropped: at 2 ms it cost 0.2 ms per step at
    N=2, where the host-side launch path is the bottleneck (profiles/r1h_dp2_ab.txt)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc, self.index = [], None, index
        self.t_begin = self.t_end = None

    def _phys_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[self.index])
            except (ValueError, IndexError):
                pass
        return self.index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self._phys_index()}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            def read():
                for l in self.proc.stdout:
                    self.lines.append((time.perf_counter(), l))
            self.th = threading.Thread(target=read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def begin(self):
        self.t_begin = time.perf_counter()

    def end(self):
        self.t_end = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        self.th.join(timeout=2)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        def parse(rows):
            sm, mx, reasons = [], [], set()
            for _, l in rows:
                f = [x.strip() for x in l.split(",")]
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except (ValueError, IndexError):
                    continue
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            return sm, mx, reasons
        inside = [r for r in self.lines if self.t_begin is not None and self.t_end is not None and
                  self.t_begin - 0.05 <= r[0] <= self.t_end + 0.05]
        sm, mx, reasons = parse(inside)
        window = "timed region (+-50 ms)"
        if not sm:
            sm, mx, reasons = parse(self.lines)
            window = "warm-up + timed region"
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window,
                "source": "nvidia-smi -lms 50 (separate process)"}


def gpu_local_cpus(phys_index):
    """CPUs of the NUMA node the GPU hangs off (NVML), intersected with this process's affinity; None when that is
    everything we may run on anyway.  Pinned staging buffers are allocated on the node of the allocating thread: a
    loader thread on the far socket costs ~20 % of the host->device rate."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(phys_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, 16)
        cpus = {i * 64 + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cur = os.sched_getaffinity(0)
        local = cpus & cur
        return local if local and local != cur else None
    except Exception:
        return None


def cpu_reference(steps, warmup, batch, dropout=True, workload="C1"):
    """The reference's CPU path: TensorFlow is not installable here, so this is the oracle port (torch eager,
    op for op: R gather+matmul calls, 15-step Python loop, dense CE, per-variable Adam) on all host threads."""
    import torch
    from oracle import keras_semantics as K
    pkg = importlib.import_module("masters-thesis_b200")
    syn = pkg.synthetic
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = K.Config(masked_sparse_ce=True, plain_attention=True) if workload == "C2" else K.Config()
    if workload == "C4":
        cfg = K.Config(R=360, n_enc=2)
    cfg = cfg if dropout else cfg.no_dropout()
    g, _ = syn.select_groups(cfg.G, remove=None) if workload == "C4" else syn.select_groups(cfg.G)
    groups = [g] * cfg.n_enc
    P = K.init_params(cfg, groups, seed=0)
    opt = K.new_opt_state(P)
    (x, cap, a0, c0), tgt = syn.make_batch(batch)
    b = (x, cap, a0, c0, tgt)
    for i in range(warmup):
        K.train_step(cfg, P, opt, groups, b, seed=i)
    t0 = time.perf_counter()
    for i in range(steps):
        K.train_step(cfg, P, opt, groups, b, seed=warmup + i)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps * 1e3, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, cores = cpu_reference(args.steps, args.warmup, args.batch * (2 if args.workload == "C4" else 1),
                                   not args.no_dropout, args.workload)
    sample = f"{args.steps} train steps of one {args.batch}-sample batch after {args.warmup} warm-up"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.workload), "global_batch": args.batch * (2 if args.workload == "C4" else 1),
                       "parallelism": "cpu", "dropout": "off" if args.no_dropout else "config.yaml rates, shared Philox contract"},
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample,
                             "note": "CPU restatement (torch eager) of the TF/Keras path, not TF (TensorFlow is not installable here)"},
            "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def stage(msg):
    """Progress line on stderr (rank-tagged): where a multi-GPU run is when something stalls."""
    print(f"[bench r{os.environ.get('RANK', '0')} {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    wd = int(os.environ.get("TNT_BENCH_WATCHDOG", "0"))
    if wd > 0:                      # dump every thread's Python stack and exit if the run takes longer than this
        import faulthandler
        faulthandler.dump_traceback_later(wd, exit=True)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=90))   # a hang aborts fast
    pkg = importlib.import_module("masters-thesis_b200")
    syn = pkg.synthetic
    ms2 = args.workload == "C4"
    B, N, T, V, U = args.batch * (2 if ms2 else 1), syn.N_VERTICES, 15, 5001, 512
    # main.py:115 -> R = 345;  main_ms.py:139 keeps all 360 parcels
    groups, out_groups = syn.select_groups(32, remove=None) if ms2 else syn.select_groups(32)
    rates = (0.0, 0.0, 0.0, 0.0, 0.0, 0.0) if args.no_dropout else (0.0, 0.2, 0.2, 0.2, 0.2, 0.2)   # config.yaml:36-41
    def new_model(bmax):
        m = pkg.NIC((groups, out_groups), U, 512, 512, 32, V, T, *rates, 0.01, 0.001, 0.00003, 0.00001,
                    batch_max=bmax, seed=42, device=dev, n_enc=2 if ms2 else 1, masked_sparse_ce=args.workload == "C2",
                    plain_attention=args.workload == "C2")
        m.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), None)   # main.py:97
        m.build(N, bmax)
        return m
    stage("building the model")
    model = new_model(B)
    n_params = model.count_params()
    NB = 4                                                          # rotating input batches (> L2 together)
    gen = torch.Generator(device=dev); gen.manual_seed(42 + rank)
    betas_d = [torch.randn(B, N, device=dev, generator=gen) for _ in range(NB)]
    caps = syn.make_captions(NB * B, T, V, seed=42 + rank).reshape(NB, B, T)
    tgts = np.stack([syn.make_target_ids(c) for c in caps])
    cap_d = [torch.from_numpy(c).to(dev) for c in caps]
    tgt_d = [torch.from_numpy(t).to(dev) for t in tgts]
    z_d = torch.zeros(B, U, device=dev)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def step_resident(i):
        j = i % NB
        model.train_step_async(betas_d[j], cap_d[j], tgt_d[j], z_d, z_d)     # ONE C call per step on any number of GPUs

    # ---- value: device-resident inputs -----------------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    stage("warm-up (device-resident steps)")
    for i in range(args.warmup):
        step_resident(i)
    sync_all()
    stage("timed region: value")
    sampler.begin()
    l0 = model.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step_resident(args.warmup + i)
    e1.record()
    sync_all()
    sampler.end()
    launches = model.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    ms_total = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- e2e: reference-facing API with pinned HOST buffers, one-hot target, metrics read back -------
    stage("e2e: host buffers through NIC.prefetch / train_step")
    aff_old, aff_local = None, gpu_local_cpus(sampler._phys_index())
    if aff_local:
        try:
            aff_old = os.sched_getaffinity(0)
            os.sched_setaffinity(0, aff_local)          # the loader side of the loop: GPU-local NUMA node
        except OSError:
            aff_old = None
    betas_h = [b.cpu().pin_memory() for b in betas_d[:2]]
    cap_h = [torch.from_numpy(c).pin_memory() for c in caps[:2]]
    # target = the generator's id array BEFORE to_categorical (data_generator_guse.py:160-162); the (B,T,V) one-hot of
    # :163 is accepted too (collapsed to ids on the device) but is 19 MB of PCIe traffic that carries 3.8 KB of content
    tg_h = [torch.from_numpy(np.ascontiguousarray(t, dtype=np.int32)).pin_memory() for t in tgts[:2]]
    z_h = torch.zeros(B, U).pin_memory()
    data_h = [((betas_h[j], cap_h[j], z_h, z_h), tg_h[j]) for j in range(2)]
    def step_e2e(i, more):
        # what NIC.fit does per batch: start the upload of batch i+1, then run batch i (metrics read back)
        if more:
            model.prefetch(data_h[(i + 1) % 2])
        return model.train_step(data_h[i % 2])
    model.prefetch(data_h[0])
    for i in range(args.warmup):
        step_e2e(i, True)
    sync_all()
    e0.record()
    for i in range(args.steps):
        last = step_e2e(args.warmup + i, i + 1 < args.steps)
    e1.record()
    sync_all()
    ms_e2e = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())
    e2e_val = world * B * args.steps / (ms_e2e * 1e-3)
    if aff_old is not None:
        os.sched_setaffinity(0, aff_old)
    h2d = betas_h[0].numel() * 4 + cap_h[0].numel() * 4 + tg_h[0].numel() * 4 + 2 * z_h.numel() * 4
    d2h = 16 * 8

    # ---- in-situ kernel timing: CUDA events on the launching stream around every kernel group of the step,
    #      inside the normal back-to-back step loop (tnt_trace).  Every rank runs these steps (they contain the
    #      data-parallel collectives); rank 0 reports.
    stage("in-situ kernel trace")
    model.trace(True)
    n_tr = 8
    for i in range(n_tr):
        step_resident(i)
    torch.cuda.synchronize(dev)
    recs = model.trace_read()
    model.trace(False)
    sync_all()
    # ---- SURVEY 8(e) "Decode": rank-sharded greedy decode, no collective on the data path (N > 1 only) ----------------
    greedy_sharded = None
    if world > 1:
        stage("rank-sharded greedy decode")
        try:
            Bs = 2048
            gs = new_model(Bs)
            model._drain()
            gs._w.copy_(model._w); gs._state.copy_(model._state)
            xs_ = torch.randn(Bs, N, device=dev, generator=gen)
            zs_ = torch.zeros(Bs, U, device=dev)
            st_s = torch.full((Bs,), syn.START_ID, dtype=torch.int32, device=dev)
            gfs = lambda: gs.greedy_predict(xs_, zs_, zs_, st_s, T, U, None, return_raw=False, return_attention=False)
            gfs(); sync_all()
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record()
            for _ in range(3):
                gfs()
            eb.record(); sync_all()
            tms = torch.tensor([ea.elapsed_time(eb) / 3], device=dev, dtype=torch.float64)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            greedy_sharded = {"value": world * Bs / (float(tms.item()) * 1e-3), "unit": "captions/s", "batch_per_gpu": Bs,
                              "n_gpus": world, "ms_per_batch": float(tms.item()), "collectives_on_data_path": 0}
            del gs, xs_, zs_
            torch.cuda.empty_cache()
        except Exception as ex:
            greedy_sharded = {"error": repr(ex)}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    stage("rank 0: phases, decode, dataset, C4, CPU baseline")
    # ---- phase timing (rank 0, N=1 semantics: these calls contain no collective) ---------------------
    hbm_peak, peak_src = peaks()
    phases = {}
    lib, ctx, st = model._libh, model._ctx, model._stream()
    xb, xc, tg = betas_d[0], cap_d[0], tgt_d[0]
    def timed(fn, reps=10):
        fn(); torch.cuda.synchronize(dev)
        a, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tot = 0.0
        for r in range(reps):
            flush.zero_()                                            # > L2: cold start for every repetition
            a.record(); fn(); b2.record(); torch.cuda.synchronize(dev)
            tot += a.elapsed_time(b2)
        return tot / reps
    flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
    if world == 1:      # after tnt_dp_init the phases carry the data-parallel exchanges: rank 0 alone must not call them
        phases["encode_ms"] = timed(lambda: lib.tnt_phase_encode(ctx, xb.data_ptr(), B, B, 1, 1, st))
        phases["decode_fwd_bwd_ms"] = timed(lambda: lib.tnt_phase_decode(ctx, xc.data_ptr(), tg.data_ptr(), z_d.data_ptr(),
                                                                          z_d.data_ptr(), B, B, 1, 1, 1, st))
        phases["encoder_bwd_ms"] = timed(lambda: lib.tnt_phase_encoder_backward(ctx, xb.data_ptr(), B, B, st))
        phases["optimizer_ms"] = timed(lambda: lib.tnt_phase_optimizer(ctx, st))
    # algorithmic bytes / flops per launch as DESIGN.md §4
    tf_peak = tensor_peak()
    agg = {}
    for name, ms in recs:
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += ms
    n_lab = int(sum(len(g) for g in groups))
    n_enc_params = int(model._bucket(0).numel())
    TB, G, E, H, R = T * B, 32, 512, 256, len(groups)
    Bh = B // 2 if ms2 else B                    # rows one encoder launch handles
    alg = {   # name prefix -> (bound, bytes or flops per launch, note)
        "enc_fwd_": ("hbm", 4.0 * (Bh * n_lab + n_lab * G), "x gathered once (B*N_lab*4) + packed kernel (N_lab*G*4)"),
        "enc_transpose": ("hbm", 8.0 * Bh * N, "x read once, xT written once"),
        "enc_bwd_dw_": ("hbm", 4.0 * (Bh * n_lab + n_lab * G), "x gathered again + dW written"),
        "opt_norm_kernel": ("hbm", 8.0 * n_params, "reads w, g"),
        "opt_adam_kernel": ("hbm", 28.0 * n_params, "reads w,m,v,g writes w,m,v"),
        "opt_norm_kernel (encoder part)": ("hbm", 8.0 * n_enc_params, "reads w, g of the encoder variables"),
        "opt_adam_kernel (encoder part)": ("hbm", 28.0 * n_enc_params, "reads w,m,v,g writes w,m,v (encoder variables)"),
        "opt_norm_kernel (decoder part)": ("hbm", 8.0 * (n_params - n_enc_params), "reads w, g of the decoder variables"),
        "opt_adam_kernel (decoder part)": ("hbm", 28.0 * (n_params - n_enc_params), "reads w,m,v,g writes w,m,v (decoder variables)"),
        "pd_fwd_kernel": ("fp32-fma", 2.0 * B * (G + U) * 4 * U * T, "gate GEMM flops of the 15 dependent steps; latency bound"),
        "pd_bwd_kernel": ("fp32-fma", 2.0 * B * (G + U) * 4 * U * T, "d x GEMM flops of the 15 dependent steps; latency bound"),
        "head forward": ("tensor", 2.0 * TB * (U * H + H * V), "dense_inter + dense_out"),
        "head backward (critical)": ("tensor", 2.0 * TB * (H * V + U * H), "d inter + d hout"),
        "head weight gradients": ("tensor", 2.0 * TB * (H * V + U * H), "d oW + d iW"),
        "decoder weight gradients": ("tensor", 2.0 * TB * (U * 4 * U + G * 4 * U + 2 * E * 4 * U + U * 32), "dU,dWc,dWe,dW2,dxtext"),
        "prologue": ("tensor", 2.0 * TB * E * 4 * U, "Zx"),
    }
    kernels = []
    for name, (cnt, tot) in agg.items():
        ms = tot / cnt
        ent = {"name": name, "ms": ms, "launches_per_step": cnt / n_tr}
        for pre, (bound, work, note) in alg.items():
            if name.startswith(pre):
                ent["bound"] = bound
                if bound == "hbm":
                    ach = work / (ms * 1e-3) / 1e9
                    ent.update({"achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "alg_bytes_per_launch": work})
                else:
                    ach = work / (ms * 1e-3) / 1e12
                    pk = tf_peak if bound == "tensor" else FP32_FMA_PEAK_TFLOPS
                    ent.update({"achieved": ach, "peak": pk, "unit": "TFLOP/s (fp32-equivalent)", "frac": ach / pk, "alg_flops_per_launch": work})
                ent["note"] = note
        kernels.append(ent)
    kernels.sort(key=lambda e: -e["ms"])
    # dominant kernel = the longest kernel group on the MAIN stream (the step's critical path; side-stream groups run
    # next to it and are listed in `kernels` too)
    main_k = [k for k in kernels if "side stream" not in k["name"]]
    dominant = main_k[0] if main_k else None
    hbm_k = [k for k in kernels if k.get("bound") == "hbm"]
    top_hbm = max(hbm_k, key=lambda k: k["alg_bytes_per_launch"]) if hbm_k else None
    alg_sample = ALG_BYTES_PER_SAMPLE_C4 if ms2 else ALG_BYTES_PER_SAMPLE
    step_achieved = alg_sample * (value / world) / 1e9
    # the path's roofline is HBM; achieved = algorithmic bytes per sample x samples/s per GPU
    roofline = {"bound": "hbm", "achieved": step_achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": step_achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                "scope": "whole train step: SURVEY 8(d) algorithmic bytes per sample x samples/s per GPU",
                "alg_bytes_per_sample": alg_sample, "alg_bytes_per_step": alg_sample * B,
                "dominant_kernel": ({"kernel": dominant["name"], "ms_per_launch": dominant["ms"],
                                     "bound": dominant.get("bound"), "achieved": dominant.get("achieved"),
                                     "peak": dominant.get("peak"), "unit": dominant.get("unit"),
                                     "frac": dominant.get("frac"), "share_of_step": dominant["ms"] / ms_step,
                                     "note": dominant.get("note")} if dominant else None),
                "largest_hbm_kernel": ({"kernel": top_hbm["name"], "achieved": top_hbm["achieved"], "frac": top_hbm["frac"],
                                        "alg_bytes_per_launch": top_hbm["alg_bytes_per_launch"],
                                        "ms_per_launch": top_hbm["ms"],
                                        "traffic": next((NCU_TRAFFIC[k] for k in sorted(NCU_TRAFFIC, key=len, reverse=True)
                                                         if top_hbm["name"].startswith(k)), None) if not ms2 else None}
                                       if top_hbm else None),
                "timing": "value: CUDA events around the K-step loop; kernels: CUDA events on the launching stream "
                          "inside the back-to-back step loop (tnt_trace)"}
    step_roofline = {"bound": "hbm", "achieved": step_achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": step_achieved / hbm_peak, "alg_bytes_per_sample": alg_sample}
    kernel_variants = model.kernel_variants()
    n_lab_bytes = 4.0 * n_lab                       # algorithmic bytes per decoded caption (ids only)

    def decode_roofline(cps):
        ach = n_lab_bytes * cps / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                "alg_bytes_per_caption": n_lab_bytes}

    # ---- greedy captions/s (config[4]): 515-caption test split, ids only -------------------------------
    greedy = None
    try:
        Bd = 515
        gm = new_model(Bd)
        model._drain()                                            # a deferred decoder update may be on the side stream
        gm._w.copy_(model._w); gm._state.copy_(model._state)
        xg = torch.randn(Bd, N, device=dev, generator=gen)
        zg = torch.zeros(Bd, U, device=dev)
        start = torch.full((Bd,), syn.START_ID, dtype=torch.int32, device=dev)
        gfn = lambda: gm.greedy_predict(xg, zg, zg, start, T, U, None, return_raw=False, return_attention=False)
        gfn(); torch.cuda.synchronize(dev)
        reps = 5
        e0.record()
        for _ in range(reps):
            gfn()
        e1.record(); torch.cuda.synchronize(dev)
        gms = e0.elapsed_time(e1) / reps
        greedy = {"value": Bd / (gms * 1e-3), "unit": "captions/s", "batch": Bd, "max_len": T, "ms_per_batch": gms,
                  "outputs": "token ids (device-resident betas, ids read back)",
                  "roofline": decode_roofline(Bd / (gms * 1e-3)), "kernel_variants": gm.kernel_variants()}
    except Exception as ex:   # decode is a secondary line; never lose the headline because of it
        greedy = {"error": repr(ex)}
    # the same decode at a large batch (configs[4]: "at large batch, inference throughput path"); world == 1 only
    greedy_large = None
    if world == 1 and not ms2 and greedy and "error" not in greedy:
        try:
            del gm, xg, zg
            torch.cuda.empty_cache()
            Bl = 4096
            gl = new_model(Bl)
            model._drain()
            gl._w.copy_(model._w); gl._state.copy_(model._state)
            xl = torch.randn(Bl, N, device=dev, generator=gen)
            zl = torch.zeros(Bl, U, device=dev)
            startl = torch.full((Bl,), syn.START_ID, dtype=torch.int32, device=dev)
            gfl = lambda: gl.greedy_predict(xl, zl, zl, startl, T, U, None, return_raw=False, return_attention=False)
            gfl(); torch.cuda.synchronize(dev)
            reps = 3
            e0.record()
            for _ in range(reps):
                gfl()
            e1.record(); torch.cuda.synchronize(dev)
            gms = e0.elapsed_time(e1) / reps
            greedy_large = {"value": Bl / (gms * 1e-3), "unit": "captions/s", "batch": Bl, "max_len": T, "ms_per_batch": gms,
                            "outputs": "token ids (device-resident betas, ids read back)",
                            "roofline": decode_roofline(Bl / (gms * 1e-3))}
            # SURVEY C5's largest size: 16 384 captions, decoded as four 4096-row chunks of one greedy_predict call
            try:
                del xl, zl
                Bx = 16384
                xx = torch.randn(Bx, N, device=dev, generator=gen)
                zx = torch.zeros(Bx, U, device=dev)
                startx = torch.full((Bx,), syn.START_ID, dtype=torch.int32, device=dev)
                gfx = lambda: gl.greedy_predict(xx, zx, zx, startx, T, U, None, return_raw=False, return_attention=False)
                gfx(); torch.cuda.synchronize(dev)
                e0.record(); gfx(); e1.record(); torch.cuda.synchronize(dev)
                gms = e0.elapsed_time(e1)
                greedy_large["b16384"] = {"value": Bx / (gms * 1e-3), "unit": "captions/s", "batch": Bx, "chunk": Bl,
                                          "ms_per_batch": gms, "roofline": decode_roofline(Bx / (gms * 1e-3))}
                del xx, zx
            except Exception as ex:
                greedy_large["b16384"] = {"error": repr(ex)}
            del gl
            torch.cuda.empty_cache()
        except Exception as ex:
            greedy_large = {"error": repr(ex)}

    # ---- batches assembled on the device from HBM-resident tables (no PCIe traffic per step) -----
    device_dataset = None
    if world == 1 and not ms2:
        try:
            n_keys, n_pairs = 512, 64 * B                       # 671 MB beta table (> L2), 64 batches per epoch
            table = torch.randn(n_keys, N, device=dev, generator=gen)
            rs = np.random.RandomState(7)
            dgen = pkg.DeviceDataGenerator(table, rs.randint(0, n_keys, n_pairs),
                                           list(syn.make_captions(n_pairs, T, V, seed=9)), B, U, T, V,
                                           shuffle=True, training=True, device=dev, rng=rs)
            for i in range(3):
                dgen[i]
            torch.cuda.synchronize(dev)
            reps = 20
            e0.record()
            for i in range(reps):
                dgen[3 + i]
            e1.record(); torch.cuda.synchronize(dev)
            asm_ms = e0.elapsed_time(e1) / reps
            asm_bytes = 2.0 * B * N * 4 + 3.0 * B * T * 4
            for i in range(args.warmup):
                model.train_step(dgen[i])
            torch.cuda.synchronize(dev)
            e0.record()
            for i in range(args.steps):
                model.train_step(dgen[(args.warmup + i) % len(dgen)])       # batch assembly + step + metrics read back
            e1.record(); torch.cuda.synchronize(dev)
            fit_ms = e0.elapsed_time(e1) / args.steps
            dgen.check()
            device_dataset = {"value": B / (fit_ms * 1e-3), "unit": "samples/s", "ms_per_step": fit_ms,
                              "api": "NIC.train_step(DeviceDataGenerator[i]): tnt_assemble_betas + tnt_assemble_captions "
                                     "from a 512-row beta table in HBM, then the step, metrics read back every step",
                              "assemble": {"ms_per_batch": asm_ms, "alg_bytes_per_batch": asm_bytes, "bound": "hbm",
                                           "achieved": asm_bytes / (asm_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                           "frac": asm_bytes / (asm_ms * 1e-3) / 1e9 / hbm_peak}}
            del dgen, table
        except Exception as ex:
            device_dataset = {"error": repr(ex)}

    # ---- configs[3]: the two-subject model, one N=1 measurement beside the headline (full line: --workload C4) -------
    c4 = None
    if world == 1 and not ms2 and args.workload == "C1":
        try:
            g4, og4 = syn.select_groups(32, remove=None)
            m4 = pkg.NIC((g4, og4), U, 512, 512, 32, V, T, *rates, 0.01, 0.001, 0.00003, 0.00001, batch_max=2 * B,
                         seed=42, device=dev, n_enc=2)
            m4.compile(pkg.Adam(1e-4, beta_1=0.9, beta_2=0.98, epsilon=10.0e-9, clipnorm=0.1), None)
            m4.build(N, 2 * B)
            xb4 = [torch.cat([betas_d[j], betas_d[(j + 1) % NB]]) for j in range(2)]
            cp4 = [torch.cat([cap_d[j], cap_d[(j + 1) % NB]]) for j in range(2)]
            tg4 = [torch.cat([tgt_d[j], tgt_d[(j + 1) % NB]]) for j in range(2)]
            z4 = torch.zeros(2 * B, U, device=dev)
            for i in range(5):
                m4.train_step_async(xb4[i % 2], cp4[i % 2], tg4[i % 2], z4, z4)
            torch.cuda.synchronize(dev)
            n4 = 50
            e0.record()
            for i in range(n4):
                m4.train_step_async(xb4[i % 2], cp4[i % 2], tg4[i % 2], z4, z4)
            e1.record(); torch.cuda.synchronize(dev)
            ms4 = e0.elapsed_time(e1) / n4
            v4 = 2 * B / (ms4 * 1e-3)
            a4 = ALG_BYTES_PER_SAMPLE_C4 * v4 / 1e9
            c4 = {"workload": WORKLOAD_C4, "value": v4, "unit": "samples/s", "ms_per_step": ms4, "steps": n4,
                  "n_params": m4.count_params(), "kernel_variants": m4.kernel_variants(),
                  "roofline": {"bound": "hbm", "achieved": a4, "peak": hbm_peak, "unit": "GB/s", "frac": a4 / hbm_peak,
                               "alg_bytes_per_sample": ALG_BYTES_PER_SAMPLE_C4}}
            del m4, xb4, cp4, tg4
            torch.cuda.empty_cache()
        except Exception as ex:
            c4 = {"error": repr(ex)}

    cpu = None
    if not args.no_cpu_baseline and world == 1:
        steps_cpu = 5
        v, ms, cores = cpu_reference(steps_cpu, 1, B, not args.no_dropout, args.workload)
        cpu = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "ms_per_step": ms,
               "sample": f"{steps_cpu} train steps of one {B}-sample batch after 1 warm-up (oracle port, torch eager)"}

    line = {"metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.workload), "global_batch": B * world, "parallelism": f"dp{world}",
                       "dropout": "off" if args.no_dropout else "config.yaml rates, in-kernel Philox",
                       "kernel_variants": kernel_variants, "precision": "fp32 (TNT_FP32: bf16x6 operand split, fp32 accumulate)",
                       "l2_policy": f"{NB} rotating resident input batches (4x84 MB) + 253 MB of w/m/v/g per step: "
                                    "working set larger than the 126 MB L2"},
            "e2e": {"value": e2e_val, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "numa_bound_cpus": (len(aff_local) if aff_old is not None else None), "api": "NIC.prefetch(batch i+1); NIC.train_step(((betas,cap,a0,c0), target_ids)) -- the loop of NIC.fit; every batch is copied from pinned host memory inside the timed region; target = the generator's (B,T) id array before to_categorical",
                    "last_metrics": last},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "step_roofline": step_roofline,
            "phases_ms": phases, "kernels": kernels, "dominant_kernel": dominant, "greedy": greedy, "greedy_large": greedy_large, "device_dataset": device_dataset, "c4": c4, "greedy_sharded": greedy_sharded, "cpu_baseline": cpu, "n_params": n_params}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
