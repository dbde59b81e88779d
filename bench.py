#!/usr/bin/env python
"""bench.py -- ODEConvGRU training step (BASELINE.json configs[1]: forward + backward + Adam, rk4 fixed-step,
batch 128 per GPU, 10 in / 10 out 64x64x1 synthetic Moving-MNIST-shaped frames) in predicted frames/s.

  python bench.py --gpus N --steps K --warmup W           (N > 1: launched by torch.distributed.run, one rank/GPU)
  python bench.py --impl reference ...                    (the reference's CPU path: reference wiring restated in
                                                           oracle/refmodel.py + restated torchdiffeq, all host threads)

One JSON line on rank 0.  `value` = frames/s with inputs resident in HBM; `e2e` = same through the public module API
with HOST (pinned) frames copied in and the loss read back every step; `roofline` = the dominant kernel (tcgen05
3x3 conv) against the measured bf16 peak / 2 (TF32) ; `cpu_baseline` = the oracle timed on this box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_IN = T_OUT = 10
RES = 64
FEVAL_FLOPS = 94371840            # per sample, 5 x conv3x3 64->64 at 16x16
CONV3_FLOPS = 2 * 256 * 64 * 576  # one 3x3 64->64 conv, per sample


NET_TRAFFIC_F16_B128 = 10748416  # bytes per launch at batch 128: dram read 9.21 MB + written 1.01-2.06 MB (mean of the two launches captured; ncu --set full, profiles/r02c_ncu.md)
NET_TRAFFIC_B128 = 10654848  # bytes per launch at batch 128: dram read 9.95 MB + written 0.42-0.98 MB (mean of the two launches captured; ncu --set full, profiles/r01_final_ncu.md)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="odex", choices=["odex", "reference"])
    ap.add_argument("--batch", type=int, default=128, help="per-GPU batch (configs[1]: 128)")
    ap.add_argument("--method", default="rk4")
    ap.add_argument("--cpu-batch", type=int, default=0,
                    help="batch of the CPU arm: 0 = the same batch as the GPU arm for --impl reference (same config), 16 for the "
                         "bounded cpu_baseline sample inside the default line")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the sub-records of the default line (dopri5 configs[2], small_batch configs[0], K4/K5 GB/s, strong scaling)")
    ap.add_argument("--no-graph", action="store_true", help="enqueue every step from Python instead of replaying a CUDA graph")
    ap.add_argument("--forward-only", action="store_true",
                    help="time the model forward alone (no_grad) instead of the training step: SURVEY.md 8d cfg3 (dopri5, batch 512)")
    ap.add_argument("--workload", default="odeconvgru", choices=["odeconvgru", "vidode"],
                    help="odeconvgru = BASELINE configs[1] (the default, the driver's line); vidode = configs[4]: VidODE forward at "
                         "128x128x3, dopri5, 32 sequences per GPU (use --batch 32)")
    return ap.parse_args()


def synth(batch, seed, device="cpu", pin=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(batch, T_IN, 1, RES, RES, generator=g)
    y = torch.rand(batch, T_OUT, 1, RES, RES, generator=torch.Generator().manual_seed(seed + 1))
    if pin:
        x, y = x.pin_memory(), y.pin_memory()
    return x.to(device), y.to(device)


def times(device):
    ts = torch.tensor(np.arange(0, T_IN + T_OUT) / (T_IN + T_OUT))
    return ts[:T_IN].to(device), ts[T_IN:].to(device)


# ------------------------------------------------------------------------------------------- CPU reference arm
def cpu_training_step_fn(method, batch):
    """fwd + MSE + bwd + Adam of the reference ODEConvGRU wiring on CPU fp32 (oracle/refmodel.py over a state_dict;
    anomaly mode off -- the reference turns it on at train_test.py:5)."""
    from ode_rl_b200.harness import ODEConvGRUHarness
    from oracle import refmodel

    torch.manual_seed(0)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in ODEConvGRUHarness(method).state_dict().items()
          if not (k.startswith("ode_convgru_cell.ode_func.") or k.startswith("diffeq_solver.ode_func."))}
    params = list(sd.values())
    full = dict(sd)
    for k in list(sd):
        if k.startswith("ode_encoder_func."):
            full["ode_convgru_cell.ode_func." + k[len("ode_encoder_func."):]] = sd[k]
        if k.startswith("ode_decoder_func."):
            full["diffeq_solver.ode_func." + k[len("ode_decoder_func."):]] = sd[k]
    opt = torch.optim.Adam(params, lr=1e-3)
    x, y = synth(batch, 1234)
    obs, tp = times("cpu")

    def step():
        opt.zero_grad(set_to_none=True)
        pred = refmodel.odeconvgru_forward(full, x, obs, tp, method)
        loss = torch.nn.functional.mse_loss(pred, y)
        loss.backward()
        opt.step()
        return float(loss)

    return step


def time_cpu(method, batch, steps, warmup):
    torch.set_num_threads(os.cpu_count() or 1)
    step = cpu_training_step_fn(method, batch)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return batch * T_OUT / dt, dt


def workload_name(method, forward_only, batch):
    """config.workload, identical for the GPU arm and the --impl reference arm."""
    cfg = ("1" if method != "dopri5" else "2 (as a training step)") if not forward_only else ("2" if method == "dopri5" else "1 (forward only)")
    what = "forward (no_grad)" if forward_only else "forward+backward+Adam"
    m = method + (" fixed-step" if method != "dopri5" else " adaptive rtol=1e-3 atol=1e-4")
    return "BASELINE configs[%s]: ODEConvGRU %s, %s, batch %d per GPU, 10 in / 10 out 64x64x1" % (cfg, what, m, batch)


def print_line(text):
    print(text, flush=True)


# ------------------------------------------------------------------------------------------- clocks sampler
class Clocks:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def summary(self, t0, t1):
        if self.p:
            self.p.terminate()
        sm, mx, reasons = [], 0, set()
        for ts, line in self.rows:
            if not (t0 <= ts <= t1 + 0.2):
                continue
            f = [s.strip() for s in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------- configs[4]: VidODE forward
def vidode_cpu(batch, steps):
    """oracle/refmodel.vidode_forward (the reference's models/VidODE.py wiring restated, pinned by tests/golden/vidode_*.pt)
    on the host cores: forward, dopri5 with the reference's default tolerances."""
    from ode_rl_b200.harness import VidODEHarness
    from oracle import refmodel

    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    sd = {k: v.detach() for k, v in VidODEHarness("dopri5").state_dict().items()}
    x = torch.rand(batch, T_IN, 3, 128, 128, generator=torch.Generator().manual_seed(1234))
    obs, tp = times("cpu")
    with torch.no_grad():
        refmodel.vidode_forward(sd, x[:1], obs, tp, "dopri5")
        t0 = time.perf_counter()
        for _ in range(steps):
            refmodel.vidode_forward(sd, x, obs, tp, "dopri5")
    dt = (time.perf_counter() - t0) / steps
    return batch * T_OUT / dt, dt


def vidode_main(args, world, rank, local):
    """BASELINE configs[4] / SURVEY.md 8d cfg5: VidODE (models/VidODE.py) forward at 128x128x3 -> 128 latent channels on
    32x32, dopri5 (the reference's DiffEqSolver defaults rtol 1e-4 / atol 1e-5), 32 sequences per GPU (256 on 8 GPUs), masks
    all ones.  predicted frames/s = B * 10 / forward time.  The hot path (ODEConvGRUCell + DiffEqSolver) runs on libodex;
    BN encoder / per-step decoder / grid_sample warp are torch -- their share is reported."""
    import torch.distributed as dist

    from ode_rl_b200 import _lib, kernels as K, parallel
    from ode_rl_b200.harness import VidODEHarness

    assert torch.cuda.is_available(), "bench.py (odex arm) needs a CUDA device; there is no CPU fallback"
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    torch.manual_seed(0)
    model = VidODEHarness("dopri5", device=dev).to(dev)
    model.device = dev
    if world > 1:
        parallel.shard_solver(model.diffeq_solver, mode="sum")     # batch-shared step size = the single-process sequence
    xh = torch.rand(B, T_IN, 3, 128, 128, generator=torch.Generator().manual_seed(1234 + rank)).pin_memory()
    x = xh.to(dev)
    obs, tp = times(dev)
    batch = {"observed_tp": obs, "tp_to_predict": tp, "observed_mask": torch.ones(B, T_IN, 1, device=dev),
             "mask_predicted_data": torch.ones(B, T_OUT, 1, device=dev)}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / 1e3, w0, time.perf_counter()

    def fwd():
        with torch.no_grad():
            return model(x, batch)[0]

    def hot():
        with torch.no_grad():
            return model.latents(x, batch)[1]

    # e2e input pipeline, as in the training bench: while forward k computes, forward k+1's frames travel host -> device on a
    # copy stream into a staging buffer; every timed forward contains one full H2D transfer of its 63 MB of frames (overlapping
    # the previous forward), one device-to-device move into the input buffer and a 4-byte synchronising read of the result
    xd, xs = torch.empty_like(x), torch.empty_like(x)
    copy_stream = torch.cuda.Stream(device=dev)

    def prefetch():
        copy_stream.wait_stream(torch.cuda.current_stream())      # the staging buffer has been consumed
        with torch.cuda.stream(copy_stream):
            xs.copy_(xh, non_blocking=True)

    def e2e_step():
        torch.cuda.current_stream().wait_stream(copy_stream)      # this forward's frames have arrived
        xd.copy_(xs, non_blocking=True)
        prefetch()
        with torch.no_grad():
            return float(model(xd, batch)[0].mean())        # 4-byte result read back (synchronising)

    for _ in range(max(3, args.warmup)):
        fwd()
    clocks = Clocks(local) if rank == 0 else None
    time.sleep(0.3)
    n0 = _lib.launch_count()
    sec, w0, w1 = timed(fwd, args.steps)
    launches = _lib.launch_count() - n0
    stats = model.diffeq_solver.last_stats
    sec_hot, _, _ = timed(hot, args.steps)
    prefetch()
    e2e_step()
    sec_e, _, w1 = timed(e2e_step, args.steps)
    clk = clocks.summary(w0, w1) if clocks else None
    value = world * B * T_OUT * args.steps / sec

    roof = cpu = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        # dominant kernel: the ConvGRU gates convolution 5x5, cat(x, h) = 256 -> 256 channels on 32x32 (conv_tc_kernel<128>, two
        # N-blocks), 10 launches per forward; repeated from C between two events on the launching stream
        be = K.backend()
        xs, hs = torch.randn(B, 32, 32, 128, device=dev), torch.randn(B, 32, 32, 128, device=dev)
        wg = model.encoder_z0.cgru_cell.conv_gates[0]
        wt = be.pack_weights(wg.weight, False, True)
        g = torch.empty(B, 32, 32, 256, device=dev)
        reps = 20
        be.conv2d(xs, hs, None, wt, 256, 5, wg.bias, K.ACT_NONE, g, repeat=3)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        be.conv2d(xs, hs, None, wt, 256, 5, wg.bias, K.ACT_NONE, g, repeat=reps)
        e1.record()
        torch.cuda.synchronize()
        avg = e0.elapsed_time(e1) / 1e3 / reps
        fl = 2 * B * 32 * 32 * 256 * 256 * 25
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        roof = {"bound": "tensor", "kernel": "conv_tc_kernel<128> (ConvGRU gates 5x5 256->256 on 32x32, 3xTF32, tcgen05 kind::tf32)",
                "achieved": fl / avg / 1e12, "peak": peak, "unit": "TFLOP/s", "frac": fl / avg / 1e12 / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of the two N-block launches of one such convolution at batch 32
                # (ncu --set full, profiles/r01_vidode_ncu.md: 40.15 MB + 0.35 MB each; the second block's inputs hit L2)
                "traffic": int(2 * (40151040 + 349184) * B / 32), "traffic_source": "profiles/r01_vidode_ncu.md (ncu --set full)",
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks else "fallback 1.4 PF (of fallback)",
                "note": "algorithmic FLOPs counted once; 3 TF32 MMAs per useful one at half the bf16 rate: frac*6 = share of tensor "
                        "issue slots against the SUSTAINED bf16 figure (it can exceed 1: a lone kernel runs at the burst clock, "
                        "ceiling = bf16 burst / 6); ncu: tensor pipe active 77 %% of SM-active cycles; %d convolutions (2 N-block "
                        "launches each) back to back, batch %d" % (reps, B),
                "launches_timed": reps, "avg_launch_us": avg * 1e6}
        if not args.no_cpu_baseline and world == 1:
            v, dt = vidode_cpu(2, 1)
            cpu = {"value": v, "unit": "frames/s", "cores": os.cpu_count() or 1, "kind": "port",
                   "sample": "1 timed forward at batch 2 (of the batch-%d workload) on the host cores; oracle/refmodel.vidode_forward + "
                             "restated torchdiffeq, torch CPU fp32" % B}
        line = {"metric": "VidODE forward predicted frames/s", "value": value, "unit": "frames/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32 (3xTF32 tensor-core convs, fp32 accumulate)", "data": "synthetic",
                "config": {"workload": "BASELINE configs[4]: VidODE forward, 128x128x3, 10 in / 10 out, dopri5 rtol=1e-4 atol=1e-5 "
                                       "(DiffEqSolver defaults, models/VidODE.py:86), batch %d per GPU, masks all ones, BatchNorm with "
                                       "per-rank batch statistics" % B,
                           "global_batch": B * world,
                           "parallelism": "batch-sharded dp%d; dopri5 error norm all-reduced (SUM) so every rank takes the global-batch "
                                          "step sequence" % world,
                           "l2": "no explicit flush: per-forward working set (frames 63 MB + latents + decoder activations) >> 126 MB L2"},
                "dopri5": {"n_accept": stats.n_accept, "n_reject": stats.n_reject, "n_feval": stats.n_feval},
                "hot_path": {"ms_per_step": sec_hot / args.steps * 1e3, "share_of_forward": sec_hot / sec,
                             "what": "conv encoder (torch) + ODEConvGRUCell + DiffEqSolver (libodex); the rest of the forward is the "
                                     "torch-side per-step decoder, grid_sample warp and composition (SURVEY.md 8f n3)"},
                "e2e": {"value": world * B * T_OUT * args.steps / sec_e, "unit": "frames/s", "h2d_bytes_per_step": int(xh.numel() * 4),
                        "d2h_bytes_per_step": 4, "ms_per_step": sec_e / args.steps * 1e3,
                        "pipeline": "pinned host -> staging buffer on a copy stream, overlapping the previous forward; staging -> "
                                    "input buffer device-to-device; result mean read back (synchronising) every forward"},
                "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "cpu_baseline": cpu}
        print_line(json.dumps(line))
    if world > 1:
        _teardown(dist, [])



# ------------------------------------------------------------------------------------------- sub-records of the default line
def _ev_time(fn, k, warm=2):
    """seconds per call of fn over k calls, CUDA events on the current stream, after `warm` untimed calls."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 1e3 / k


def sub_dopri5(dev, B=512):
    """BASELINE configs[2]: ODEConvGRU dopri5 (rtol 1e-3 / atol 1e-4), batch 512, device-resident step control, batch-shared
    step size: forward (no_grad) and training step, with the solver's counters; default-init weights and the x3 'stiff-ish'
    variant of SURVEY.md 8d in which attempts get rejected."""
    from ode_rl_b200.harness import ODEConvGRUHarness

    out = {"batch": B, "rtol": 1e-3, "atol": 1e-4}
    x, y = synth(B, 4321, dev)
    obs, tp = times(dev)
    for name, scale in (("default_init", 1.0), ("stiff_x3", 3.0)):
        torch.manual_seed(0)
        model = ODEConvGRUHarness("dopri5", rtol=1e-3, atol=1e-4, device=dev).to(dev)
        if scale != 1.0:
            with torch.no_grad():
                for p_ in model.ode_decoder_func.parameters():
                    p_.mul_(scale)
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)

        def fwd():
            with torch.no_grad():
                return model(x, obs, tp, return_latents=False)

        def train():
            opt.zero_grad(set_to_none=True)
            loss = torch.nn.functional.mse_loss(model(x, obs, tp, return_latents=False), y)
            loss.backward()
            opt.step()

        t_f = _ev_time(fwd, 5)
        st = model.diffeq_solver.last_stats
        rec = {"forward_ms": t_f * 1e3, "forward_frames_per_s": B * T_OUT / t_f, "n_accept": st.n_accept, "n_reject": st.n_reject,
               "n_feval": st.n_feval, "host_reads_per_forward": st.host_syncs}
        if scale == 1.0:
            opt_state = None
            t_t = _ev_time(train, 3)
            rec.update({"training_ms": t_t * 1e3, "training_frames_per_s": B * T_OUT / t_t})
        out[name] = rec
        del model, opt
    torch.cuda.empty_cache()
    return out


def sub_small_batch(dev, with_cpu=True):
    """BASELINE configs[0]: ODEConvGRU forward, batch 16, euler -- the reference's own CPU-runnable case -- on the GPU (eager
    and as one replayed CUDA graph: at this size the ~260 launches, not the tensor work, decide the time) beside the CPU oracle
    at the SAME batch; and the reference's test-time rollout length (configs.yaml:628-630: up to 190 predicted frames)."""
    from ode_rl_b200 import _lib
    from ode_rl_b200.harness import ODEConvGRUHarness

    B = 16
    torch.manual_seed(0)
    model = ODEConvGRUHarness("euler", device=dev).to(dev).eval()
    x, _ = synth(B, 99, dev)
    obs, tp = times(dev)
    out = {"batch": B, "method": "euler"}

    def fwd(tp_=tp):
        with torch.no_grad():
            return model(x, obs, tp_, return_latents=False)

    n0 = _lib.launch_count()
    fwd()
    launches = _lib.launch_count() - n0
    t_e = _ev_time(fwd, 20)
    # the same forward captured once and replayed
    static_x = x.clone()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            with torch.no_grad():
                model(static_x, obs, tp, return_latents=False)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        with torch.no_grad():
            static_out = model(static_x, obs, tp, return_latents=False)
    t_g = _ev_time(g.replay, 50)
    out.update({"forward_ms_eager": t_e * 1e3, "forward_ms_graph": t_g * 1e3, "frames_per_s": B * T_OUT / t_g,
                "libodex_launches_per_forward": int(launches), "launches_per_predicted_frame": launches / T_OUT})
    # long rollout: 10 observed frames -> 190 predicted (rk4), batch 16
    model.diffeq_solver.ode_method = "rk4"
    ts = torch.tensor(np.arange(0, 200) / 200.0)
    tp_long = ts[10:].to(dev)
    n0 = _lib.launch_count()
    t_l = _ev_time(lambda: fwd(tp_long), 3, warm=1)
    n_l = (_lib.launch_count() - n0) / 4
    out["long_rollout"] = {"predicted_frames": 190, "method": "rk4", "forward_ms": t_l * 1e3, "frames_per_s": B * 190 / t_l,
                           "libodex_launches_per_predicted_frame": n_l / 190}
    model.diffeq_solver.ode_method = "euler"
    if with_cpu:
        from oracle import refmodel
        torch.set_num_threads(os.cpu_count() or 1)
        sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        xc = x.cpu()
        oc, tc = times("cpu")
        with torch.no_grad():
            ref = refmodel.odeconvgru_forward(sd, xc, oc, tc, "euler")
            t0 = time.perf_counter()
            for _ in range(3):
                refmodel.odeconvgru_forward(sd, xc, oc, tc, "euler")
            t_c = (time.perf_counter() - t0) / 3
        err = float((static_out.cpu() - ref).abs().max())
        out.update({"cpu_forward_ms": t_c * 1e3, "cpu_frames_per_s": B * T_OUT / t_c, "cpu_cores": os.cpu_count() or 1,
                    "gpu_over_cpu": t_c / t_g, "max_abs_diff_vs_cpu_oracle": err})
    del model
    return out


def sub_vidode(dev, B=32):
    """BASELINE configs[4] inside the default line: VidODE forward (128x128x3, dopri5 at the reference's rtol 1e-4 / atol 1e-5)
    at the per-GPU batch of the 8-GPU configuration (256 / 8 = 32), device-resident frames.  The full record (e2e, hot-path share,
    roofline of its top kernel, CPU arm) is `bench.py --workload vidode`."""
    from ode_rl_b200.harness import VidODEHarness
    torch.manual_seed(0)
    model = VidODEHarness("dopri5", device=dev).to(dev)
    model.device = dev
    x = torch.rand(B, T_IN, 3, 128, 128, generator=torch.Generator().manual_seed(1234)).to(dev)
    obs, tp = times(dev)
    batch = {"observed_tp": obs, "tp_to_predict": tp, "observed_mask": torch.ones(B, T_IN, 1, device=dev),
             "mask_predicted_data": torch.ones(B, T_OUT, 1, device=dev)}

    def fwd():
        with torch.no_grad():
            return model(x, batch)[0]

    sec = _ev_time(fwd, 5, warm=3)
    st = model.diffeq_solver.last_stats
    out = {"workload": "cfg5 VidODE dopri5 forward, 128x128x3, batch %d (per-GPU share of batch 256 on 8 GPUs)" % B,
           "value": B * T_OUT / sec, "unit": "frames/s", "ms_per_forward": sec * 1e3}
    if st is not None:
        out.update(n_accept=int(st.n_accept), n_reject=int(st.n_reject), n_feval=int(st.n_feval))
    del model, x
    torch.cuda.empty_cache()
    return out


def sub_rk_kernels(dev, be, hbm, batches=(512, 2048)):
    """Achieved GB/s (algorithmic bytes of SURVEY.md 8d / kernel time) of the dopri5 kernels north_star names: K4 =
    dopri_errnorm_kernel (error estimate + RMS norm + accept/reject + controller in one launch: 8 planes read = 32 B/elem) and
    K5 = dopri_advance_kernel (dense output for n_out output times + FSAL advance: 8 reads + n_out writes for the outputs, plus
    the y / k0 advance writes, 2 more planes)."""
    from ode_rl_b200 import _lib

    res = {}
    for nB in batches:
        planes = [torch.randn(nB, 16, 16, 64, device=dev) * 0.1 for _ in range(9)]
        y0, y1, k7 = planes[0], planes[1], planes[2:9]
        n_el = y0.numel()
        st = torch.zeros(be.dopri_state_bytes(), dtype=torch.uint8, device=dev)
        t_dev = torch.tensor([0.0, 0.01, 1e300], dtype=torch.float64, device=dev)     # the last output is never reached
        sol = torch.empty((3,) + tuple(y0.shape), device=dev)
        log = torch.zeros((4096, 4), dtype=torch.float64, device=dev)
        st64 = st[:_lib.STATE_HEAD_BYTES].view(torch.float64)

        def reset():
            be.dopri_init(st, t_dev, 3, 1e9, 1e9, 1 << 30, 0)        # huge tolerances: every attempt is accepted
            st64[_lib.OFF_DT // 8] = 1e-3
            st64[_lib.OFF_DT_TRY // 8] = 1e-3

        reps = 20
        reset()
        for i in range(3):
            be.dopri_errctrl(st, t_dev, y0, y1, k7, log, 4096, i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reset()
        e0.record()
        for i in range(reps):
            be.dopri_errctrl(st, t_dev, y0, y1, k7, log, 4096, i)
        e1.record()
        torch.cuda.synchronize()
        t4 = e0.elapsed_time(e1) / 1e3 / reps
        # K5: every launch finds its own sequence number published -> runs the full dense-output + advance pass (1 output inside)
        st32 = st[:_lib.STATE_HEAD_BYTES].view(torch.int32)
        adv, jlo, jhi = _lib.DopriState.adv_seq.offset // 4, _lib.DopriState.jlo.offset // 4, _lib.DopriState.jhi.offset // 4
        reset()
        st64[_lib.DopriState.dt_acc.offset // 8] = 1e-3
        st64[_lib.DopriState.t1.offset // 8] = 0.02
        st32[adv], st32[jlo], st32[jhi] = 7, 1, 2
        for _ in range(3):
            be.dopri_advance(st, t_dev, y0, y1, k7, sol, None, None, 0, 7)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            be.dopri_advance(st, t_dev, y0, y1, k7, sol, None, None, 0, 7)
        e1.record()
        torch.cuda.synchronize()
        t5 = e0.elapsed_time(e1) / 1e3 / reps
        b4, b5 = 8 * n_el * 4, (8 + 1 + 2) * n_el * 4
        res["B%d" % nB] = {"plane_mb": n_el * 4 / 1e6,
                           "K4_errnorm_ctrl": {"us": t4 * 1e6, "bytes": b4, "GBps": b4 / t4 / 1e9, "frac": b4 / t4 / 1e9 / hbm},
                           "K5_dense_advance": {"us": t5 * 1e6, "bytes": b5, "GBps": b5 / t5 / 1e9, "frac": b5 / t5 / 1e9 / hbm,
                                                "outputs_in_step": 1}}
        del planes, sol
    torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------------------- GPU arm
def main():
    args = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and args.impl != "reference":
        # stdout carries exactly ONE line (the JSON): NCCL prints its version banner to stdout from native code on the first
        # collective, so file descriptor 1 points at stderr until the line is printed
        sys.stdout.flush()
        _saved_out = os.dup(1)
        os.dup2(2, 1)
        global print_line
        print_line = lambda text: (sys.stdout.flush(), os.dup2(_saved_out, 1), print(text, flush=True))
    if args.workload == "vidode" and args.impl != "reference":
        return vidode_main(args, world, rank, local)

    if args.impl == "reference" and args.workload == "vidode":
        if rank != 0:
            return
        cb = min(args.cpu_batch or 4, 4)
        val, dt = vidode_cpu(cb, max(1, args.steps))
        cores = os.cpu_count() or 1
        print(json.dumps({
            "impl": "reference", "metric": "VidODE forward predicted frames/s", "value": val, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "VidODE forward, 128x128x3, 10 in / 10 out, dopri5 rtol=1e-4 atol=1e-5", "batch_per_step": cb},
            "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port",
                             "sample": "batch %d forwards of the configs[4] workload (models/VidODE.py wiring restated in "
                                       "oracle/refmodel.vidode_forward + restated torchdiffeq 0.2.1; torch CPU, %d threads)" % (cb, cores)},
            "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    if args.impl == "reference":
        if rank != 0:
            return
        cpu_batch = args.cpu_batch or args.batch
        # bounded: at batch 128 one CPU training step is seconds; at most 4 timed steps + 1 warm-up (the GPU arm's K / W are sized
        # for a 20 ms step)
        k_cpu = max(1, min(args.steps, 4))
        val, dt = time_cpu(args.method, cpu_batch, k_cpu, 1)
        args.cpu_batch = cpu_batch
        cores = os.cpu_count() or 1
        line = {"impl": "reference", "metric": "ODEConvGRU training-step predicted frames/s", "value": val,
                "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload_name(args.method, False, args.cpu_batch), "global_batch": args.cpu_batch,
                           "batch_per_step": args.cpu_batch, "timed_steps": k_cpu,
                           "parallelism": "one process, torch CPU with %d threads" % cores},
                "cpu_baseline": {"value": val, "unit": "frames/s", "cores": cores, "kind": "port",
                                 "sample": "batch %d training steps of the configs[1] workload (reference wiring restated in "
                                           "oracle/refmodel.py + restated torchdiffeq 0.2.1; torch CPU, %d threads, anomaly "
                                           "mode off)" % (args.cpu_batch, cores)},
                "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import torch.distributed as dist

    from ode_rl_b200 import _lib, kernels as K
    from ode_rl_b200.harness import ODEConvGRUHarness

    assert torch.cuda.is_available(), "bench.py (odex arm) needs a CUDA device; there is no CPU fallback"
    torch.backends.cudnn.allow_tf32 = False          # torch-side ConvEncode/ConvDecode stay fp32 like the CPU reference
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    torch.manual_seed(0)
    model = ODEConvGRUHarness(args.method, rtol=1e-3, atol=1e-4, device=dev).to(dev)
    params = [p for p in model.parameters()]
    use_graph = (not args.no_graph) and args.method != "dopri5" and not args.forward_only   # dopri5 reads its controller state on the host
    # torch's single-kernel Adam (same update rule as the reference's torch.optim.Adam, train_test.py; 20.31 -> 20.14 ms/step);
    # ODEX_FUSED_ADAM=0 selects the default foreach implementation
    fused_adam = os.environ.get("ODEX_FUSED_ADAM", "1") == "1"
    opt = torch.optim.Adam(params, lr=1e-3, capturable=use_graph, fused=fused_adam) if fused_adam else \
        torch.optim.Adam(params, lr=1e-3, capturable=use_graph)
    from ode_rl_b200 import parallel
    if world > 1:
        parallel.shard_solver(model.diffeq_solver, mode="sum")                # dopri5: batch-shared step size (SUM = parity mode)
    # ODEX_OVERLAP_AR=1: gradient buckets in the order the backward finishes them -- decoder side (ConvDecode + the solver's ODE
    # function), encoder cell, then ConvEncode -- the first two all-reduced from hooks on a side stream, under the rest of the
    # backward.  Captures and runs (2 x B200: 13.47 ms against 13.42 ms with one bucket after the backward): the collective is not
    # what the multi-GPU step waits for, so the default stays one bucket.
    overlap_ar = world > 1 and os.environ.get("ODEX_OVERLAP_AR", "0") == "1" and not args.forward_only
    stages = [list(model.conv_decoder.parameters()) + list(model.ode_decoder_func.parameters()),
              list(model.ode_convgru_cell.parameters())] if overlap_ar else None
    bucket = parallel.GradBucket(params, stages=stages)
    xh, yh = synth(B, 1234 + rank, "cpu", pin=True)
    x, y = xh.to(dev), yh.to(dev)
    obs, tp = times(dev)

    def step_eager(xd, yd):
        if args.forward_only:
            with torch.no_grad():
                return torch.nn.functional.mse_loss(model(xd, obs, tp, return_latents=False), yd)
        opt.zero_grad(set_to_none=True)
        pred = model(xd, obs, tp, return_latents=False)
        loss = torch.nn.functional.mse_loss(pred, yd)
        loss.backward()
        if world > 1:
            bucket.allreduce()
        opt.step()
        return loss

    graphed = None
    if use_graph:
        # the whole step (zero_grad, forward, loss, backward, Adam; the NCCL all-reduce stays eager between the captured
        # backward and the optimizer) is captured once and replayed: ~1200 launches/step no longer pass through Python
        from ode_rl_b200.harness import GraphedTrainStep
        graphed = GraphedTrainStep(model, opt, torch.nn.functional.mse_loss, x, y, obs, tp,
                                   grad_hook=(bucket.allreduce if world > 1 else None), warmup=3)

    def step(xd, yd):
        if graphed is None:
            return step_eager(xd, yd)
        if xd is x:
            return graphed()               # inputs already resident in the graph's static buffers
        return graphed(xd, yd)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / 1e3, w0, time.perf_counter()

    for _ in range(max(3, args.warmup)):
        step(x, y)
    clocks = Clocks(local) if rank == 0 else None
    time.sleep(0.3)
    n0 = _lib.launch_count()
    sec, w0, w1 = timed(lambda: step(x, y), args.steps)
    launches = _lib.launch_count() - n0
    if graphed is not None:
        launches = graphed.launches_per_step * args.steps      # kernels inside the replayed graph
    value = world * B * T_OUT * args.steps / sec

    # ---- e2e: host (pinned) frames -> device every step, loss read back every step
    xd, yd = torch.empty_like(x), torch.empty_like(y)

    # Input pipeline of a training loop: while step k computes, step k+1's frames travel host -> device on a copy stream into
    # a staging buffer; at the start of a step the staged frames are moved into the step's input buffers (device to device)
    # and the next transfer is started.  Every timed step therefore contains one full H2D transfer of one step's inputs
    # (h2d_bytes_per_step) and one D2H read of its loss; the transfer overlaps the previous step's kernels.
    copy_stream = torch.cuda.Stream(device=dev)
    xs, ys = torch.empty_like(x), torch.empty_like(y)

    def prefetch():
        cur = torch.cuda.current_stream()
        copy_stream.wait_stream(cur)               # the staging buffers have been consumed
        with torch.cuda.stream(copy_stream):
            xs.copy_(xh, non_blocking=True)
            ys.copy_(yh, non_blocking=True)

    def e2e_step():
        torch.cuda.current_stream().wait_stream(copy_stream)        # this step's frames have arrived
        if graphed is not None:
            graphed.x.copy_(xs, non_blocking=True)
            graphed.y.copy_(ys, non_blocking=True)
            prefetch()
            return float(graphed())                # replay, loss read-back (4 bytes, synchronising)
        xd.copy_(xs, non_blocking=True)
        yd.copy_(ys, non_blocking=True)
        prefetch()
        return float(step(xd, yd).detach())

    prefetch()
    e2e_step()
    sec_e, _, w1 = timed(e2e_step, args.steps)
    e2e_value = world * B * T_OUT * args.steps / sec_e
    clk = clocks.summary(w0, w1) if clocks else None

    # ---- roofline of the dominant kernel family (tcgen05 3x3 conv 64->64: forward and dgrad of every f-eval).
    # The launch is repeated from C between two CUDA events on the launching stream (a Python loop over ctypes is slower
    # than the kernel); inputs are one state plane (B x 64 KiB), resident in L2 as they are inside the step, where every
    # layer reads what the previous kernel just wrote.
    be = K.backend()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    roof = None
    rk = None
    if rank == 0:
        # the dominant kernel of the step: the fused ODE-function chain (5 x conv3x3 64->64, ReLU between, saved activations
        # stored, `y + dt*f` residual on the last layer) = one f-eval of the training forward, 92 launches per step
        # (forward + data-gradient chains).  Launched back to back from Python: at ~66 us per launch the enqueue keeps ahead.
        L_net = 5
        xs = torch.randn(B, 16, 16, 64, device=dev)
        from ode_rl_b200 import ops as _ops
        split = 1 if _ops.CONV_SPLIT == "f16" else 0
        wts = [be.pack_weights(torch.randn(64, 64, 3, 3, device=dev) * 0.05, False, 1 + split) for _ in range(L_net)]
        bss = [torch.randn(64, device=dev) for _ in range(L_net)]
        outs_ = [torch.empty(B, 16, 16, 64, device=dev) for _ in range(L_net)]
        acts_ = [K.ACT_RELU] * (L_net - 1) + [K.ACT_NONE]
        aux_ = [None] * (L_net - 1) + [xs]
        kinds_ = [0] * (L_net - 1) + [3]
        reps = 100

        def net_pass(n):
            for _ in range(n):
                be.convnet(xs, wts, bss, acts_, aux_, kinds_, outs_, alpha=0.05, split=split)

        net_pass(10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        net_pass(reps)
        e1.record()
        torch.cuda.synchronize()
        avg = e0.elapsed_time(e1) / 1e3 / reps
        ach = L_net * CONV3_FLOPS * B / avg / 1e12
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        kname = ("convnet_f16_kernel (5 x conv3x3 64->64 fused, two-term fp16 operand split, tcgen05 kind::f16)" if split else
                 "convnet_tc_kernel (5 x conv3x3 64->64 fused, 3xTF32, tcgen05 kind::tf32)")
        roof = {"bound": "tensor", "kernel": kname, "achieved": ach,
                "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch at batch 128 (ncu --set full,
                # profiles/r01_final_ncu.md), scaled with the batch; None until that capture exists
                "traffic": (NET_TRAFFIC_F16_B128 if split else NET_TRAFFIC_B128) and int((NET_TRAFFIC_F16_B128 if split else NET_TRAFFIC_B128) * B / 128),
                "traffic_source": "profiles/r02c_ncu.md (ncu --set full, per launch)" if split else "profiles/r01_final_ncu.md (ncu --set full, per launch)",
                "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (of measured)" if peaks else "fallback 1.4 PF (of fallback)",
                "note": ("achieved counts algorithmic FLOPs once (2*M*N*K per layer, 5 layers per launch); fp32-level accuracy costs 3 fp16 "
                         "MMAs per useful one (hi.hi, hi.lo, lo.hi) at the bf16 / fp16 rate, so the ceiling of useful FLOPs is 1/3 of "
                         "the peak and frac*3 is the share of tensor issue slots used; " if split else
                         "achieved counts algorithmic FLOPs once (2*M*N*K per layer, 5 layers per launch); the kernel issues 3 TF32 "
                         "MMAs per useful one and TF32 runs at half the bf16 rate, so frac*6 is the share of tensor issue slots used; ")
                        + "%d launches back to back, batch %d" % (reps, B),
                "launches_timed": reps, "avg_launch_us": avg * 1e6}

        # ---- RK-stage GB/s (the second half of BASELINE's metric): the fused stage-combination / error-norm / dense-output
        # kernels at the dopri5 configuration's size (B=512: 33.5 MB planes, 10 planes = 335 MB > the 126 MB L2).
        nB = 512
        planes = [torch.randn(nB, 16, 16, 64, device=dev) for _ in range(9)]
        outp = torch.empty_like(planes[0])
        n_el = planes[0].numel()
        dt_dev = torch.tensor([0.05], dtype=torch.float64, device=dev)
        from ode_rl_b200.solver import DP_BETA

        def rk_pass():
            # the six stage combinations of one dopri5 attempt (26 reads + 6 writes of a plane) ...
            for s_ in range(6):
                coefs = [c for c in DP_BETA[s_] if c != 0.0]
                ins = [planes[0]] + [planes[1 + j] for j, c in enumerate(DP_BETA[s_]) if c != 0.0]
                be.lincomb(outp, ins, [1.0] + coefs, dtmask=(1 << len(ins)) - 2, dt_dev=dt_dev)
            # ... and the rk4 (3/8 rule) combinations of one step (14 reads + 4 writes)
            be.lincomb(outp, planes[:2], [1.0, 1.0 / 3.0], dtmask=2, dt_dev=dt_dev)
            be.lincomb(outp, planes[:3], [1.0, -1.0 / 3.0, 1.0], dtmask=6, dt_dev=dt_dev)
            be.lincomb(outp, planes[:4], [1.0, 1.0, -1.0, 1.0], dtmask=14, dt_dev=dt_dev)
            be.lincomb(outp, planes[:5], [1.0, 0.125, 0.375, 0.375, 0.125], dtmask=30, dt_dev=dt_dev)

        rk_bytes = (26 + 6 + 14 + 4) * n_el * 4
        for _ in range(3):
            rk_pass()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            rk_pass()
        e1.record()
        torch.cuda.synchronize()
        rk_s = e0.elapsed_time(e1) / 1e3 / 10
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        rk = {"bound": "hbm", "kernel": "lincomb_kernel<NIN> (dopri5 stage combines of one attempt + rk4 3/8-rule combines of one step)",
              "achieved": rk_bytes / rk_s / 1e9, "peak": hbm, "unit": "GB/s", "frac": rk_bytes / rk_s / 1e9 / hbm,
              "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
              "bytes_per_pass": rk_bytes, "planes_mb": n_el * 4 / 1e6, "batch": nB,
              "note": "algorithmic bytes: every distinct operand plane read once, every result written once (SURVEY.md 8d)"}
        del planes, outp

    # ---- strong scaling: the global batch fixed at 512, 512 / N sequences per GPU
    strong = None
    if args.method != "dopri5" and not args.forward_only and not args.no_extras and 512 % world == 0:
        try:
            Bs = 512 // world
            xs_, ys_ = synth(Bs, 777 + rank, dev)
            from ode_rl_b200.harness import GraphedTrainStep
            torch.manual_seed(0)
            m2 = ODEConvGRUHarness(args.method, rtol=1e-3, atol=1e-4, device=dev).to(dev)
            p2 = [p_ for p_ in m2.parameters()]
            o2 = torch.optim.Adam(p2, lr=1e-3, capturable=True, fused=True)
            b2 = parallel.GradBucket(p2)
            g2 = GraphedTrainStep(m2, o2, torch.nn.functional.mse_loss, xs_, ys_, obs, tp,
                                  grad_hook=(b2.allreduce if world > 1 else None), warmup=3)
            for _ in range(3):
                g2()
            sec_s, _, _ = timed(lambda: g2(), max(4, args.steps // 2))
            k_s = max(4, args.steps // 2)
            strong = {"global_batch": 512, "batch_per_gpu": Bs, "ms_per_step": sec_s / k_s * 1e3,
                      "value": 512 * T_OUT * k_s / sec_s, "unit": "frames/s", "scaling": "strong"}
            torch.cuda.synchronize()
            g2.graph.reset()
            del g2, m2, o2, b2
            torch.cuda.empty_cache()
        except Exception as e:                       # a sub-record must never cost the line
            strong = {"error": repr(e)[:200]}

    extras = {}
    if rank == 0 and not args.no_extras and args.method != "dopri5" and not args.forward_only:
        for name, fn in (("dopri5", lambda: sub_dopri5(dev)),
                         ("small_batch", lambda: sub_small_batch(dev, with_cpu=not args.no_cpu_baseline and world == 1)),
                         ("rk_kernels", lambda: sub_rk_kernels(dev, be, float(peaks.get("hbm_gbs", 6650.0)))),
                         ("vidode", lambda: sub_vidode(dev))):
            try:
                extras[name] = fn()
            except Exception as e:
                extras[name] = {"error": repr(e)[:300]}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            cb = args.cpu_batch or 16
            v, dt = time_cpu(args.method, cb, 2, 1)
            cpu = {"value": v, "unit": "frames/s", "cores": os.cpu_count() or 1, "kind": "port",
                   "sample": "2 timed training steps at batch %d (of the batch-%d workload) on the host cores; oracle/refmodel.py + "
                             "restated torchdiffeq, torch CPU fp32, anomaly mode off" % (cb, B)}
        line = {"metric": "ODEConvGRU %s predicted frames/s" % ("forward" if args.forward_only else "training-step"), "value": value,
                "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": sec / args.steps * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (tensor-core convs from split operands: two-term fp16 on kind::f16 / 3xTF32 on kind::tf32; fp32 accumulate)",
                "data": "synthetic",
                "config": {"workload": workload_name(args.method, args.forward_only, B),
                           "global_batch": B * world, "parallelism": "batch-sharded dp%d, NCCL flat-bucket gradient allreduce" % world,
                           "l2": "no explicit flush: the per-step working set (saved activations ~2 GB) is >> the 126 MB L2",
                           "cuda_graph": graphed is not None,
                           "allreduce_in_graph": bool(graphed is not None and world > 1 and graphed.hook_in_graph),
                           "allreduce_overlapped": bool(overlap_ar),
                           "optimizer": "torch.optim.Adam(fused=%s)" % fused_adam},
                "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": int(xh.numel() * 4 + yh.numel() * 4),
                        "d2h_bytes_per_step": 4, "ms_per_step": sec_e / args.steps * 1e3,
                        "pipeline": "pinned host -> staging buffer on a copy stream, overlapping the previous step; staging -> step "
                                    "inputs device-to-device at the start of the step; loss read back (synchronising) every step"},
                "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "rk_stage": rk, "cpu_baseline": cpu,
                "strong_scaling": strong}
        if rk is not None and "rk_kernels" in extras:
            rk["K4_K5"] = extras.pop("rk_kernels")
        line.update(extras)
        print_line(json.dumps(line))
    if world > 1:
        _teardown(dist, [graphed])


def _teardown(dist, graphs):
    """Leave a multi-rank run without hanging: a CUDA graph that holds captured NCCL work must be released before the process
    group is destroyed, and a stuck communicator teardown (seen once with the all-reduce captured in the step graph: every rank
    had printed / finished, destroy_process_group never returned) must not keep the launcher alive -- the line is already out,
    so a watchdog ends the process with status 0 if teardown takes longer than 20 s."""
    import gc
    import threading

    sys.stdout.flush()
    sys.stderr.flush()
    t = threading.Timer(20.0, lambda: os._exit(0))
    t.daemon = True
    t.start()
    try:
        torch.cuda.synchronize()
        for g in graphs:
            if g is not None and getattr(g, "graph", None) is not None:
                g.graph.reset()
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
    finally:
        t.cancel()


if __name__ == "__main__":
    main()
