#!/usr/bin/env python
"""bench.py -- libear-b200 benchmark (contract in the task prompt / DESIGN.md section 6).

Headline metric (BASELINE.json): object gain vectors/s for extent objects on BS.2051 9+10+3
(config 1: 1M objects per GPU, P = 1,652-point spreading grid), plus the GainInterpolator
render throughput (config 2: 128 objects x 60 s x 48 kHz into 4+9+0) as a second block.

  python bench.py --gpus N --steps K --warmup W            our arm (one rank per GPU)
  python bench.py --impl reference ...                     libear's CPU path on the host cores

One "step" = one pass of the hot path over one batch of synthetic input.  `value` is
measured with inputs resident in HBM (CUDA events on the launching stream, max over
ranks); `e2e` goes through the reference-facing C-ABI host entry points with pinned HOST
buffers, host<->device copies inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LAYOUT = "9+10+3"
N_OBJECTS = 1_000_000            # per GPU (weak scaling)
FLOP_PER_PASS = 2 * 1652 * 22    # contraction flops per spreading pass
RENDER = dict(N=128, M=14, T=2_880_000, ramp=512, layout="4+9+0")


def synth_objects(n, seed):
    """SURVEY.md 8(d) cfg 2: directions uniform on the sphere, distance U(0.2, 1.5), width and
    height U(0, 360), depth 0 for half the objects and U(0, 1) for the rest, gain/diffuse U(0, 1)."""
    rng = np.random.default_rng(seed)
    az = rng.uniform(-180, 180, n)
    el = np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
    dist = rng.uniform(0.2, 1.5, n)
    w, h = rng.uniform(0, 360, n), rng.uniform(0, 360, n)
    depth = np.where(rng.uniform(size=n) < 0.5, 0.0, rng.uniform(0, 1, n))
    return [az, el, dist, w, h, depth, rng.uniform(0, 1, n), rng.uniform(0, 1, n)]


class ClockSampler:
    """SM clocks and throttle reasons sampled DURING the timed regions: an NVML polling thread
    (5 ms period) that only records between start() and stop(); samples accumulate over all
    timed regions of the run.  Falls back to one `nvidia-smi` query per window without NVML."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
               "sw_power_cap": 0x4}

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.sm, self.smax, self.reasons = [], [], set()
        self.active = False
        self.thread = None
        self.handle = None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a list of indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = gpu_index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if gpu_index < len(ids) and ids[gpu_index].isdigit():
                    phys = int(ids[gpu_index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.stop_flag = False
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.smax.append(self.max_mhz)
        try:
            mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
        except Exception:
            mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        for name, bit in self.REASONS.items():
            if mask & bit:
                self.reasons.add(name)

    def _poll(self):
        while not self.stop_flag:
            if self.active:
                try:
                    self._sample_nvml()
                except Exception:
                    pass
            time.sleep(0.005)

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            f = [x.strip() for x in out.strip().splitlines()[0].split(",")]
            self.sm.append(float(f[0]))
            self.smax.append(float(f[1]))
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[2:6]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)
        except Exception:
            pass

    def start(self):
        self.active = True
        if self.nvml is None:
            self._sample_smi()

    def stop(self):
        """Ends one sampling window; the polling thread sampled while the host sat in the closing
        synchronize, i.e. while the GPU was running the timed steps."""
        self.active = False

    def summary(self):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["no clock samples"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": float(max(self.smax)),
                "samples": len(self.sm), "reasons": sorted(self.reasons),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            return json.load(f)[kernel]["bytes"]
    except (OSError, KeyError, ValueError):
        return None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return p.get("hbm_gbs", 6650.0), p.get("bf16_tflops", 1590.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1590.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------ CPU baseline (libear itself)
def cpu_objects_throughput(args_list, threads, target_seconds):
    """libear's GainCalculatorObjects::calculate<float> on `threads` host threads, one calculator
    instance per thread (instances are not re-entrant), on a bounded sample of the workload
    (each thread walks its share of the objects, repeating it until ~target_seconds)."""
    from oracle import ref
    calcs = [ref.Objects(LAYOUT) for _ in range(threads)]
    probe = [a[:2000] for a in args_list]
    t0 = time.perf_counter()
    calcs[0].calculate(*probe, dtype=np.float32)
    per_obj = (time.perf_counter() - t0) / 2000
    share = len(args_list[0]) // threads
    per_thread = int(max(2000, min(share, target_seconds / per_obj)))
    reps = int(max(1, round(target_seconds / (per_thread * per_obj))))

    def work(i):
        sl = slice(i * per_thread, (i + 1) * per_thread)
        sub = [a[sl] for a in args_list]
        for _ in range(reps):
            calcs[i].calculate(*sub, dtype=np.float32)

    ths = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    [t.start() for t in ths]
    [t.join() for t in ths]
    dt = time.perf_counter() - t0
    n = per_thread * threads * reps
    return n / dt, n, dt


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    """--impl reference: the reference's own CPU implementation on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libear_ref.so not built"}))
        return
    threads = host_threads()
    # the GPU arm's own input: the same 10^6 objects (rank 0's seed), every one of them per step
    objs = synth_objects(args.objects, 2)
    step_n = len(objs[0])
    calcs = [ref.Objects(LAYOUT) for _ in range(threads)]
    per = -(-step_n // threads)

    def step():
        def work(i):
            sl = slice(i * per, min(step_n, (i + 1) * per))
            if sl.start < step_n:
                calcs[i].calculate(*[a[sl] for a in objs], dtype=np.float32)
        ths = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
        [t.start() for t in ths]
        [t.join() for t in ths]

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = step_n * args.steps / dt
    line = {
        "impl": "reference", "metric": "object gain vectors/s (9+10+3, extent)", "value": value,
        "unit": "vectors/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"configs[1] (SURVEY cfg 2): {step_n} extent objects/GPU on {LAYOUT} via the batched "
                               "GainCalculatorObjects C-ABI (1,652-point spreading grid, depth != 0 for half)",
                   "layout": LAYOUT, "objects_per_gpu": step_n,
                   "passes_per_step_per_gpu": step_n + int(np.count_nonzero(objs[5])),
                   "l2": "inputs+outputs 256 MB/step/GPU > 126 MB L2", "sharding": "objects, no collective"},
        "note": "libear's GainCalculatorObjects::calculate<float> on the host cores, the GPU arm's own input (same "
                "seed), all of it every step",
        "cpu_baseline": {"value": value, "unit": "vectors/s", "cores": threads, "kind": "reference",
                         "sample": f"{step_n} objects/step x {args.steps} steps, libear "
                                   f"GainCalculatorObjects::calculate<float>, xsimd ISA {ref.extent_isa()}"},
        "e2e": {"value": value, "unit": "vectors/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def ref_objects_threaded(ref, args_list, threads):
    """libear on a list of object columns, one calculator per thread (checker, outside any timing)."""
    n = len(args_list[0])
    per = -(-n // threads)
    outs = [None] * threads

    def work(i):
        sl = slice(i * per, min(n, (i + 1) * per))
        if sl.start < n:
            outs[i] = ref.Objects(LAYOUT).calculate(*[a[sl] for a in args_list], dtype=np.float32)
    ths = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    [t.start() for t in ths]
    [t.join() for t in ths]
    outs = [o for o in outs if o is not None]
    return np.concatenate([o[0] for o in outs]), np.concatenate([o[1] for o in outs])


# ------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    import libear_b200 as eb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libear-b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on stdout at the first collective; stdout carries exactly
        # one JSON line, so the banner is sent to stderr
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    hbm_peak, bf16_peak, peak_src = measured_peaks()
    stream = torch.cuda.current_stream().cuda_stream
    sampler = ClockSampler(local_rank)
    # what the timed regions produced is compared with libear (oracle/_ref, the checker) AFTER the
    # timing, on samples; a failed check makes the run exit non-zero
    parity = {"checked_against": "libear (oracle/_ref) on samples of the buffers the timed regions wrote", "ok": True}
    check = rank == 0 and not args.no_parity
    if check:
        from oracle import ref as oref
        check = oref.available()
        if not check:
            parity = {"ok": None, "note": "oracle/_ref/libear_ref.so not built: nothing checked"}

    def sustained(step_fn, units_per_step, seconds):
        """the same step back to back for `seconds` of device time (clocks settle under load)"""
        if seconds <= 0:
            return None
        s_sampler = ClockSampler(local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step_fn()
        e1.record()
        torch.cuda.synchronize()
        k = int(max(10, seconds * 1e3 / max(e0.elapsed_time(e1), 1e-3)))
        barrier()
        s_sampler.start()
        e0.record()
        for _ in range(k):
            step_fn()
        e1.record()
        barrier()
        s_sampler.stop()
        ms = max_over_ranks(e0.elapsed_time(e1))
        cl = s_sampler.summary()
        return {"value": units_per_step * world * k / (ms * 1e-3), "steps": k, "seconds": ms * 1e-3,
                "ms_per_step": ms / k, "sm_mhz": cl["sm_mhz"], "reasons": cl["reasons"]}

    # ---------------- objects: BASELINE config 1 (1M extent objects per GPU on 9+10+3)
    n = args.objects
    objs = synth_objects(n, 2 + rank)
    n_passes = n + int(np.count_nonzero(objs[5]))
    gc = eb.GainCalculatorObjects(LAYOUT, device=local_rank)
    C_out = gc.num_outputs
    d_in = torch.from_numpy(np.stack(objs)).to(dev)  # [8][n] float64
    d_direct = torch.empty((n, C_out), dtype=torch.float32, device=dev)
    d_diffuse = torch.empty((n, C_out), dtype=torch.float32, device=dev)
    soa = [d_in[k].data_ptr() for k in range(8)]

    def step_objects():
        gc.calculate_device(n, soa, d_direct.data_ptr(), d_diffuse.data_ptr(), stream=stream)

    for _ in range(args.warmup):
        step_objects()
    barrier()
    gc.set_profiling(True)
    launches0 = gc.launch_count
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step_objects()
    ev1.record()
    barrier()
    sampler.stop()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    core_ms, core_launches, _ = gc.kernel_time()
    gc.set_profiling(False)
    gpu_launches = gc.launch_count - launches0
    value = n * world * args.steps / (ms_total * 1e-3)
    core_tflops = FLOP_PER_PASS * n_passes * args.steps / (core_ms * 1e-3) / 1e12
    weight_evals = 1652.0 * n_passes * args.steps / (core_ms * 1e-3)

    # end to end through the C-ABI host entry point, pinned host buffers
    h_in = [torch.from_numpy(a).pin_memory() for a in objs]
    h_direct = torch.empty((n, C_out), dtype=torch.float32).pin_memory()
    h_diffuse = torch.empty((n, C_out), dtype=torch.float32).pin_memory()
    e2e_steps = max(1, min(args.steps, 5))

    import ctypes as C
    soa_h = eb._ObjectsSoA(*[t.data_ptr() for t in h_in], None)

    def step_e2e_abi():
        eb._chk(eb._lib.earb200_objects_calculate_host(gc.h, C.c_int64(n), C.byref(soa_h), 0,
                                                       C.c_void_p(h_direct.data_ptr()),
                                                       C.c_void_p(h_diffuse.data_ptr()), None))

    if args.no_e2e:   # profiler runs only: the line then carries no end-to-end number
        e2e_value, e2e_checksum, e2e_floor = None, None, None
    else:
        step_e2e_abi()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_e2e_abi()
        torch.cuda.synchronize()
        e2e_dt = max_over_ranks(time.perf_counter() - t0)
        e2e_value = n * world * e2e_steps / e2e_dt
        e2e_checksum = float(h_direct[:1000].double().sum())
        # the floor the host path cannot go below on this box: the same bytes, copies only (inputs in
        # on one stream while the outputs of the step go out on another), no kernels
        s_a, s_b = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        d_cols = [torch.empty(n, dtype=torch.float64, device=dev) for _ in range(8)]

        def copies():
            with torch.cuda.stream(s_a):
                for dc, hc in zip(d_cols, h_in):
                    dc.copy_(hc, non_blocking=True)
            with torch.cuda.stream(s_b):
                h_direct.copy_(d_direct, non_blocking=True)
                h_diffuse.copy_(d_diffuse, non_blocking=True)
        copies()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            copies()
        torch.cuda.synchronize()
        floor_dt = max_over_ranks(time.perf_counter() - t0)
        e2e_floor = {"value": n * world * e2e_steps / floor_dt, "unit": "vectors/s", "ms_per_step": floor_dt / e2e_steps * 1e3,
                     "what": "host<->device copies of one step alone (pinned, both directions at once), no kernels: "
                             "the PCIe / host-memory bound of the end-to-end path on this box"}
        del d_cols

    objects_sustained = sustained(step_objects, n, args.sustained)
    if check:
        # a strided sample of the rows the timed device path wrote, and the host path's rows
        sl = slice(7, n, max(1, n // 20_000))
        got_d, got_f = d_direct[sl].cpu().numpy(), d_diffuse[sl].cpu().numpy()
        exp_d, exp_f = ref_objects_threaded(oref, [a[sl] for a in objs], host_threads())
        fin = np.isfinite(exp_d).all(1) & np.isfinite(exp_f).all(1)   # libear's own acos(>1) NaNs, ~1 in 3e5
        parity["gains_max_abs"] = float(max(np.abs(got_d[fin] - exp_d[fin]).max(), np.abs(got_f[fin] - exp_f[fin]).max()))
        parity["gains_n_checked"] = int(fin.sum())
        parity["gains_bar"] = 1e-5
        if e2e_value is not None:
            parity["e2e_rows_equal_device_rows"] = bool(
                np.array_equal(h_direct[sl].numpy().view(np.uint32), got_d.view(np.uint32)) and
                np.array_equal(h_diffuse[sl].numpy().view(np.uint32), got_f.view(np.uint32)))
            parity["ok"] = parity["ok"] and parity["e2e_rows_equal_device_rows"]
        parity["ok"] = bool(parity["ok"] and parity["gains_max_abs"] < 1e-5 and fin.mean() > 0.999)

    # ---------------- render: BASELINE config 2 (128 objects x 60 s x 48 kHz into 4+9+0)
    render = None
    if not args.no_render:
        R = RENDER
        N, M, T = R["N"], R["M"], R["T"]
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        x = (torch.randn((N, T), generator=g, device=dev, dtype=torch.float32) * 0.25).clamp_(-1, 1)
        y = torch.empty((M, T), dtype=torch.float32, device=dev)
        samples = np.arange(0, T + R["ramp"], R["ramp"], dtype=np.int64)
        rng = np.random.default_rng(3 + rank)
        gains = (rng.uniform(0, 1, size=(len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
        gi = eb.GainInterpolator(N, M, device=local_rank)
        gi.set_points(samples, gains)

        from libear_b200.sharding import connect_bus
        rbus = None
        if world > 1 and args.exchange == "bus":
            # object-sharded mixing (128 objects per GPU into ONE bus): the render kernel's epilogue
            # stores every tile into the slot of the GPU that owns its time slice (peer stores over
            # NVLink); the owners sum; the summed slices are gathered into every rank's full bus
            rbus = connect_bus(eb.Bus(M, T, rank=rank, n_ranks=world, device=local_rank))

            def step_render(gather=bool(args.bus_gather)):
                gi.process_device_bus(0, T, x.data_ptr(), T, rbus, 0, stream=stream)
                if gather:
                    rbus.reduce(gather=True, stream=stream)
                    rbus.wait_result(stream=stream)
                else:
                    rbus.reduce(d_out=y.data_ptr(), out_stride=T, stream=stream)
        elif world > 1:
            # comparison arm: render, then one NCCL all-reduce of the bus
            def step_render():
                gi.process_device(0, T, x.data_ptr(), T, y.data_ptr(), T, stream=stream)
                dist.all_reduce(y)
        else:
            def step_render():
                gi.process_device(0, T, x.data_ptr(), T, y.data_ptr(), T, stream=stream)

        for _ in range(args.warmup):
            step_render()
        barrier()
        gi.set_profiling(True)
        rl0 = gi.launch_count
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        r0.record()
        for _ in range(args.steps):
            step_render()
        r1.record()
        barrier()
        sampler.stop()
        r_ms = max_over_ranks(r0.elapsed_time(r1))
        k_ms, k_launches = gi.kernel_time()
        gi.set_profiling(False)
        gpu_launches += gi.launch_count - rl0
        alg_bytes = 4.0 * T * (N + M) + 4.0 * N * M * len(samples)
        achieved = alg_bytes * args.steps / (k_ms * 1e-3) / 1e9
        rs_ms = None
        if rbus is not None and args.bus_gather:
            # the same step ending with every rank holding its time slice only (reduce-scatter)
            step_render(gather=False)
            barrier()
            r0.record()
            for _ in range(args.steps):
                step_render(gather=False)
            r1.record()
            barrier()
            rs_ms = max_over_ranks(r0.elapsed_time(r1)) / args.steps
        render = {
            "metric": "render Msamples/s (input object-samples)",
            "value": N * T * world * args.steps / (r_ms * 1e-3) / 1e6, "unit": "Msamples/s",
            "ms_per_step": r_ms / args.steps, "ms_per_step_reduce_scatter": rs_ms,
            "config": {"workload": f"configs[2] (SURVEY cfg 3): {N} objects/GPU x 60 s x 48 kHz into {R['layout']} (M={M}), "
                                   f"{R['ramp']}-sample linear gain ramps; inputs 1.6 GB/step > L2",
                       "bus_reduce": ("node-wide bus over peer memory (earb200_bus_*): tiles stored into the owners' "
                                      "slots by the render epilogue, deterministic sum per time slice"
                                      + (", gather to every rank" if args.bus_gather else ", no gather (reduce-scatter)")
                                      if rbus is not None else "nccl all_reduce(sum, f32) of the [M x T] bus after the render")
                       if world > 1 else "none"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak,
                         "traffic": ncu_traffic("render_tc_persistent_kernel") if (N, M, T) == (128, 14, 2_880_000) else None,
                         "peak_source": peak_src,
                         "kernel": "render_tc_persistent_kernel<32>", "kernel_ms": k_ms / max(1, k_launches),
                         "algorithmic_bytes_per_launch": alg_bytes},
        }
        if world == 1:
            render["sustained"] = sustained(step_render, N * T / 1e6, args.sustained)
        if check and world == 1:
            wr = np.random.default_rng(17)
            peak = float(y.abs().max())
            worst = 0.0
            starts = [0, T - 512] + [int(t) for t in wr.integers(0, T - 512, 14)]
            for t0 in starts:
                exp = oref.interp_matrix(x[:, t0:t0 + 512].cpu().numpy(), samples, gains, block_start0=t0)
                worst = max(worst, float(np.abs(y[:, t0:t0 + 512].cpu().numpy() - exp).max()) / peak)
            parity["render_audio_max_over_peak"] = worst
            parity["render_windows_checked"] = len(starts)
            parity["render_bar"] = 2e-6
            parity["ok"] = parity["ok"] and worst < 2e-6
        if world > 1 and not args.no_parity:
            # rank 0 renders the objects of ALL ranks for a window alone and compares with the summed bus
            Tw = 8192
            xw = [torch.empty((N, Tw), dtype=torch.float32, device=dev) for _ in range(world)]
            dist.all_gather(xw, x[:, :Tw].contiguous())
            gw = [None] * world
            dist.all_gather_object(gw, gains[:Tw // R["ramp"] + 1])
            step_render()
            torch.cuda.synchronize()
            if rank == 0:
                got = torch.empty((M, Tw), dtype=torch.float32, device=dev)
                if rbus is not None and args.bus_gather:
                    rbus.copy_result(got.data_ptr(), Tw, 0, Tw, stream=stream)
                else:   # the NCCL arm's bus, or rank 0's slice of the reduce-scatter (it starts at sample 0)
                    got.copy_(y[:, :Tw])
                want = torch.zeros((M, Tw), dtype=torch.float32, device=dev)
                for r in range(world):
                    gr = eb.GainInterpolator(N, M, device=local_rank)
                    gr.set_points(samples[:Tw // R["ramp"] + 1], gw[r])
                    gr.set_accumulate(r > 0)
                    gr.process_device(0, Tw, xw[r].data_ptr(), Tw, want.data_ptr(), Tw, stream=stream)
                torch.cuda.synchronize()
                err = float((got - want).abs().max() / want.abs().max())
                parity["sharded_render_vs_one_gpu_max_over_peak"] = err
                parity["ok"] = parity["ok"] and err < 1e-6
            del xw
        # end to end: pinned host input/output through the C-ABI host entry point
        if world == 1 and not args.no_render_e2e:
            hx = torch.empty((N, T), dtype=torch.float32).pin_memory()
            hx.copy_(x)
            hy = torch.empty((M, T), dtype=torch.float32).pin_memory()

            def render_e2e():
                eb._chk(eb._lib.earb200_interp_process_host(gi.h, C.c_int64(0), C.c_int64(T),
                                                            C.c_void_p(hx.data_ptr()), C.c_int64(T),
                                                            C.c_void_p(hy.data_ptr()), C.c_int64(T)))
            render_e2e()
            t0 = time.perf_counter()
            for _ in range(2):
                render_e2e()
            dt = time.perf_counter() - t0
            render["e2e"] = {"value": N * T * 2 / dt / 1e6, "unit": "Msamples/s",
                             "h2d_bytes_per_step": 4 * N * T, "d2h_bytes_per_step": 4 * M * T}
            del hx, hy
        del x, y


    # ---------------- point sources: BASELINE configs[0] (1,024 objects on 0+5+0, no extent) as the
    # reference's own CPU-runnable case -- latency-bound -- and the same kernels at 10^6 objects
    point_block = None
    if not args.no_points:
        gcp = eb.GainCalculatorObjects("0+5+0", device=local_rank)
        Cp = gcp.num_outputs
        point_block = {"metric": "point-source gain vectors/s (0+5+0, no extent): configs[0] (SURVEY cfg 1)", "unit": "vectors/s"}
        for label, npts in (("n1024", 1024), ("n1M", 1_000_000)):
            rp = np.random.default_rng(1 + rank)
            cols = [rp.uniform(-180, 180, npts), np.degrees(np.arcsin(rp.uniform(-1, 1, npts))), np.ones(npts),
                    np.zeros(npts), np.zeros(npts), np.zeros(npts), rp.uniform(0, 1, npts), rp.uniform(0, 1, npts)]
            dp_in = torch.from_numpy(np.stack(cols)).to(dev)
            dp_d = torch.empty((npts, Cp), dtype=torch.float32, device=dev)
            dp_f = torch.empty((npts, Cp), dtype=torch.float32, device=dev)
            soap = [dp_in[k].data_ptr() for k in range(8)]
            for _ in range(args.warmup):
                gcp.calculate_device(npts, soap, dp_d.data_ptr(), dp_f.data_ptr(), stream=stream)
            barrier()
            pl0 = gcp.launch_count
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record()
            for _ in range(args.steps):
                gcp.calculate_device(npts, soap, dp_d.data_ptr(), dp_f.data_ptr(), stream=stream)
            p1.record()
            barrier()
            p_ms = max_over_ranks(p0.elapsed_time(p1)) / args.steps
            gpu_launches += gcp.launch_count - pl0
            p_bytes = npts * (64.0 + 2 * 4 * Cp)
            if check and label == "n1024":
                # configs[0] is "checked against libear's CPU path": all 1,024 rows
                ed, ef = oref.Objects("0+5+0").calculate(*cols, dtype=np.float32)
                perr = float(max(np.abs(dp_d.cpu().numpy() - ed).max(), np.abs(dp_f.cpu().numpy() - ef).max()))
                parity["point_sources_1024_gains_max_abs"] = perr
                parity["ok"] = bool(parity["ok"] and perr < 1e-6)
            point_block[label] = {"objects_per_gpu": npts, "value": npts * world / (p_ms * 1e-3), "us_per_call": p_ms * 1e3,
                                  "algorithmic_gbs": p_bytes / (p_ms * 1e-3) / 1e9,
                                  "hbm_frac": p_bytes / (p_ms * 1e-3) / 1e9 / hbm_peak}
            del dp_in, dp_d, dp_f
        point_block["note"] = ("n1024 is launch-latency-bound (three small kernels per call); n1M runs the FP64 "
                               "region scan at 112 algorithmic bytes per object")
        del gcp

    # ---------------- the north star's "5200-point" wording: the extent panner on the 5,200-point
    # t-design instead of libear's 1,652-point grid.  Throughput only: libear has no such path, so
    # there is nothing to compare the gains with
    p5200_block = None
    if not args.no_points:
        n52 = min(n, 262_144)
        gc52 = eb.GainCalculatorObjects(LAYOUT, device=local_rank, grid="tdesign5200")
        d52 = torch.empty((n52, C_out), dtype=torch.float32, device=dev)
        f52 = torch.empty((n52, C_out), dtype=torch.float32, device=dev)
        for _ in range(args.warmup):
            gc52.calculate_device(n52, soa, d52.data_ptr(), f52.data_ptr(), stream=stream)
        barrier()
        q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l52 = gc52.launch_count
        q0.record()
        for _ in range(args.steps):
            gc52.calculate_device(n52, soa, d52.data_ptr(), f52.data_ptr(), stream=stream)
        q1.record()
        barrier()
        q_ms = max_over_ranks(q0.elapsed_time(q1)) / args.steps
        gpu_launches += gc52.launch_count - l52
        passes52 = n52 + int(np.count_nonzero(objs[5][:n52]))
        p5200_block = {
            "metric": "object gain vectors/s (9+10+3, extent, 5,200-point grid)", "unit": "vectors/s",
            "value": n52 * world / (q_ms * 1e-3), "ms_per_step": q_ms, "objects_per_gpu": n52,
            "contraction_tflops": 2.0 * 5200 * 22 * passes52 / (q_ms * 1e-3) / 1e12,
            "weight_evals_per_s": 5200.0 * passes52 / (q_ms * 1e-3),
            "kernel": "extent_core_tc_kernel (every weight evaluated, TF32 x 3; the cell-culled core's tables hold "
                      "2,048 points)",
            "note": "THROUGHPUT ONLY, parity unpinned: libear's extent panner has 1,652 points "
                    "(src/object_based/polar_extent.cpp:12-39, 104); the 5,200-point t-design is the grid the north "
                    "star names, and there is no reference output on it",
        }
        del gc52, d52, f52

    # ---------------- convolver: BASELINE config 4 (24 channels x 10 min, 512-tap decorrelators, B = 512)
    conv_block = None
    if not args.no_convolver:
        Cc, Bc = 24, 512
        Kc = args.conv_blocks
        Tc = Kc * Bc
        xs = (torch.randn((Cc, Tc), generator=torch.Generator(device=dev).manual_seed(200 + rank), device=dev,
                          dtype=torch.float32) * 0.25).clamp_(-1, 1)
        ys = torch.empty_like(xs)
        conv = eb.BlockConvolver(Bc, Cc, 1, device=local_rank)
        filt = eb.design_decorrelators(LAYOUT, np.float32)
        for c in range(Cc):
            fid, _ = conv.add_filter(filt[c])
            conv.set_filter(fid, c)

        def step_conv():
            conv.process_device(Kc, xs.data_ptr(), Tc, ys.data_ptr(), Tc, stream=stream)

        for _ in range(args.warmup):
            step_conv()
        barrier()
        conv.set_profiling(True)
        cl0 = conv.launch_count
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        c0.record()
        for _ in range(args.steps):
            step_conv()
        c1.record()
        barrier()
        sampler.stop()
        c_ms = max_over_ranks(c0.elapsed_time(c1))
        ck_ms, ck_n = conv.kernel_time()
        conv.set_profiling(False)
        gpu_launches += conv.launch_count - cl0
        c_bytes = 8.0 * Cc * Tc
        c_achieved = c_bytes * ck_n / (ck_ms * 1e-3) / 1e9
        conv_block = {
            "metric": "decorrelation convolver Msamples/s", "value": Cc * Tc * world * args.steps / (c_ms * 1e-3) / 1e6,
            "unit": "Msamples/s", "ms_per_step": c_ms / args.steps,
            "config": {"workload": f"configs[3] (SURVEY cfg 4): {Cc} channels/GPU x {Tc} samples ({Tc / 48000 / 60:.1f} min at 48 kHz), "
                                   f"designDecorrelators({LAYOUT}) 512-tap filters, block size {Bc}; "
                                   "5.5 GB/step > L2", "sharding": "replicated channels per GPU, no collective"},
            "roofline": {"bound": "hbm", "achieved": c_achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": c_achieved / hbm_peak,
                         "traffic": ncu_traffic("convolver_wide_kernel") if (Cc, Kc) == (24, 56250) else None,
                         "peak_source": peak_src,
                         "kernel": "convolver_wide_kernel", "kernel_ms": ck_ms / max(1, ck_n),
                         "algorithmic_bytes_per_launch": c_bytes,
                         "note": "8 B/sample; the packed-FP32 butterflies (about 3,000 warp instructions per unit "
                                 "of 3,072 samples, three quarters of them FFMA2 / FADD2 / FMUL2) bound the kernel "
                                 "before HBM does (DESIGN.md section 3, K4)"},
        }
        if world == 1:
            conv_block["sustained"] = sustained(step_conv, Cc * Tc / 1e6, args.sustained)
        if check and world == 1 and Kc > 16:
            wr = np.random.default_rng(5)
            peak = float(ys.abs().max())
            worst = 0.0
            picks = [(0, 0), (Cc - 1, Kc - 9)] + [(int(wr.integers(0, Cc)), int(wr.integers(1, Kc - 9))) for _ in range(4)]
            for c, k0 in picks:
                # one block of history: the tail of block k0 - 1; the stream continues from step to
                # step over the same input, so block 0 follows the last block of the step before
                hist = xs[c, (k0 - 1) * Bc:k0 * Bc] if k0 > 0 else xs[c, (Kc - 1) * Bc:Kc * Bc]
                rc = oref.Convolver(Bc, 1)
                fid, _ = rc.add_filter(filt[c])
                rc.set_filter(fid)
                exp = rc.process(torch.cat([hist, xs[c, k0 * Bc:(k0 + 8) * Bc]]).cpu().numpy())[Bc:]
                worst = max(worst, float(np.abs(ys[c, k0 * Bc:(k0 + 8) * Bc].cpu().numpy() - exp).max()) / peak)
            parity["convolver_audio_max_over_peak"] = worst
            parity["convolver_slices_checked"] = len(picks)
            parity["ok"] = parity["ok"] and worst < 1e-6
        if world == 1 and not args.no_render_e2e:
            hx = torch.empty((Cc, Tc), dtype=torch.float32).pin_memory()
            hx.copy_(xs)
            hy = torch.empty((Cc, Tc), dtype=torch.float32).pin_memory()

            def conv_e2e():
                eb._chk(eb._lib.earb200_convolver_process_host(conv.h, C.c_int64(Kc), C.c_void_p(hx.data_ptr()),
                                                               C.c_int64(Tc), C.c_void_p(hy.data_ptr()), C.c_int64(Tc)))
            conv_e2e()
            t0 = time.perf_counter()
            conv_e2e()
            dt = time.perf_counter() - t0
            conv_block["e2e"] = {"value": Cc * Tc / dt / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 4 * Cc * Tc,
                                 "d2h_bytes_per_step": 4 * Cc * Tc}
            del hx, hy
        del xs, ys, conv

    # ---------------- Objects signal flow: 128 objects x 60 s into 9+10+3 through
    # two interpolated renders, the decorrelator bank and the compensation delay + sum
    flow_block = None
    if not args.no_flow:
        Nf, Tf = 128, 2_880_000
        fr = eb.ObjectsRenderer(LAYOUT, Nf, device=local_rank)
        Cf = fr.num_outputs
        rngf = np.random.default_rng(7 + rank)
        f_samples = np.arange(0, Tf + 512, 512, dtype=np.int64)
        gd = (rngf.uniform(0, 1, size=(len(f_samples), Nf, Cf)) / np.sqrt(Nf)).astype(np.float32)
        gf = (rngf.uniform(0, 1, size=(len(f_samples), Nf, Cf)) / np.sqrt(Nf)).astype(np.float32)
        fr.set_gains(f_samples, gd, gf)
        del gd, gf
        xf = (torch.randn((Nf, Tf), device=dev, dtype=torch.float32) * 0.25).clamp_(-1, 1)
        yf = torch.empty((Cf, Tf), device=dev, dtype=torch.float32)
        f_stream = torch.cuda.current_stream().cuda_stream

        def step_flow():
            fr.process_device(0, Tf // 512, xf.data_ptr(), Tf, yf.data_ptr(), Tf, stream=f_stream)
        for _ in range(args.warmup):
            step_flow()
        barrier()
        fl0 = fr.launch_count
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        f0.record()
        for _ in range(args.steps):
            step_flow()
        f1.record()
        barrier()
        sampler.stop()
        f_ms = max_over_ranks(f0.elapsed_time(f1)) / args.steps
        gpu_launches += fr.launch_count - fl0
        # one pass over the objects' audio and the feeds is the least any implementation moves;
        # the pipeline as built reads the input once from HBM and once from L2 (both renders on one
        # launch) and passes two 24-channel buses through HBM
        f_bytes = 4.0 * Tf * (Nf + Cf) + 2 * 4.0 * Nf * Cf * len(f_samples)
        flow_block = {
            "metric": "Objects signal flow Msamples/s (input object-samples)",
            "value": Nf * Tf * world / (f_ms * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": f_ms,
            "config": {"workload": f"{Nf} objects/GPU x {Tf} samples into {LAYOUT} ({Cf} feeds): direct + diffuse "
                                   "gain ramps every 512 samples, designDecorrelators bank (B = 512), 255-sample "
                                   "compensation delay, sum; inputs 1.5 GB/step > L2",
                       "sharding": "objects replicated per GPU, no collective"},
            "roofline": {"bound": "hbm", "achieved": f_bytes / (f_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": f_bytes / (f_ms * 1e-3) / 1e9 / hbm_peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "render_tc_persistent_kernel (direct + diffuse paired) + convolver_fast_kernel<add> "
                                   "+ delay_advance_kernel",
                         "algorithmic_bytes_per_step": f_bytes,
                         "note": "whole pipeline against the bytes of a single fused pass"},
        }
        del xf, yf, fr

    # ---------------- HOA decode + object-sharded mixing with an NCCL bus reduce: BASELINE config 5
    mix_block = None
    if not args.no_mix:
        lay5 = "4+7+0"
        T5 = args.mix_samples
        n_total = args.mix_objects
        from libear_b200.sharding import connect_bus, object_groups, reduce_bus, shard_range
        obj_a, obj_b = shard_range(n_total, rank, world)
        n_local = obj_b - obj_a
        # HOA order 3 (ACN 0..15, SN3D): design on the device, apply as a static 16 -> 12 matrix
        hoa = eb.GainCalculatorHOA(lay5, device=local_rank)
        orders = [n for n in range(4) for _ in range(2 * n + 1)]
        degrees = [m for n in range(4) for m in range(-n, n + 1)]
        hoa.calculate(orders, degrees, "SN3D")
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        D, _ = hoa.calculate(orders, degrees, "SN3D")
        hoa_design_ms = (time.perf_counter() - t0) * 1e3
        M5 = D.shape[1]
        gi_hoa = eb.GainInterpolator(16, M5, device=local_rank)
        gi_hoa.set_points(np.array([0], dtype=np.int64), D.astype(np.float32)[None])
        xh = torch.randn((16, T5), device=dev, dtype=torch.float32) * 0.25
        yh = torch.empty((M5, T5), device=dev, dtype=torch.float32)
        for _ in range(args.warmup):
            gi_hoa.process_device(0, T5, xh.data_ptr(), T5, yh.data_ptr(), T5, stream=stream)
        barrier()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        hl0 = gi_hoa.launch_count
        h0.record()
        for _ in range(args.steps):
            gi_hoa.process_device(0, T5, xh.data_ptr(), T5, yh.data_ptr(), T5, stream=stream)
        h1.record()
        barrier()
        hoa_ms = max_over_ranks(h0.elapsed_time(h1)) / args.steps
        gpu_launches += gi_hoa.launch_count - hl0
        del xh, yh

        # objects: gains for slowly moving extent objects from the Objects calculator on the device,
        # one interpolation point per second, groups of <= 256 objects mixed into one bus
        gc5 = eb.GainCalculatorObjects(lay5, device=local_rank)
        n_pts = T5 // 48000 + 1
        samples5 = np.arange(n_pts, dtype=np.int64) * 48000
        base = synth_objects(n_local, 500 + rank)
        rng5 = np.random.default_rng(600 + rank)
        drift = rng5.uniform(-3, 3, n_local)
        gains5 = np.empty((n_pts, n_local, M5), dtype=np.float32)
        for k in range(n_pts):
            az = (base[0] + drift * k + 180.0) % 360.0 - 180.0
            d_k, _ = gc5.calculate(az, base[1], base[2], base[3], base[4], base[5], base[6], 0.0)
            gains5[k] = d_k
        groups = []
        for g0, g1 in object_groups(n_local, 256):
            gi = eb.GainInterpolator(g1 - g0, M5, device=local_rank)
            gi.set_points(samples5, gains5[:, g0:g1])
            gi.set_accumulate(g0 > 0)
            groups.append((g0, g1, gi))
        xo = torch.empty((n_local, T5), device=dev, dtype=torch.float32)
        xo.normal_(0, 0.25, generator=torch.Generator(device=dev).manual_seed(700 + rank))
        bus = torch.empty((M5, T5), device=dev, dtype=torch.float32)
        red0, red1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        mbus, slice5 = None, None
        if world > 1 and args.exchange == "bus":
            mbus = connect_bus(eb.Bus(M5, T5, rank=rank, n_ranks=world, device=local_rank))
            s0, s1 = mbus.slice()
            slice5 = torch.empty((M5, max(1, s1 - s0)), dtype=torch.float32, device=dev)

        def step_mix(timed_reduce=False, gather=bool(args.bus_gather)):
            # the groups of <= 256 objects accumulate into this rank's partial bus; with the node-wide
            # bus the LAST group's render stores the finished partial tiles into the owners' slots
            for g0, g1, gi in groups[:-1] if mbus is not None else groups:
                gi.process_device(0, T5, xo[g0].data_ptr(), T5, bus.data_ptr(), T5, stream=stream)
            if mbus is not None:
                g0, g1, gi = groups[-1]
                gi.process_device_bus(0, T5, xo[g0].data_ptr(), T5, mbus, 0, d_partial=bus.data_ptr(),
                                      partial_stride=T5, stream=stream)
            if timed_reduce:
                red0.record()
            if mbus is not None and gather:
                mbus.reduce(gather=True, stream=stream)
                mbus.wait_result(stream=stream)
            elif mbus is not None:
                mbus.reduce(d_out=slice5.data_ptr(), out_stride=slice5.shape[1], stream=stream)
            elif world > 1:
                reduce_bus(bus)
            if timed_reduce:
                red1.record()

        for _ in range(args.warmup):
            step_mix()
        barrier()
        ml0 = sum(gi.launch_count for _, _, gi in groups)
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        m0.record()
        for _ in range(args.steps):
            step_mix()
        m1.record()
        barrier()
        sampler.stop()
        mix_ms = max_over_ranks(m0.elapsed_time(m1)) / args.steps
        gpu_launches += (sum(gi.launch_count for _, _, gi in groups) - ml0)
        reduce_ms = None
        if world > 1:
            step_mix(timed_reduce=True)
            torch.cuda.synchronize()
            reduce_ms = max_over_ranks(red0.elapsed_time(red1))
        mix_rs_ms = None
        if mbus is not None and args.bus_gather:
            step_mix(gather=False)
            barrier()
            m0.record()
            for _ in range(args.steps):
                step_mix(gather=False)
            m1.record()
            barrier()
            mix_rs_ms = max_over_ranks(m0.elapsed_time(m1)) / args.steps
        mix_bytes = 4.0 * T5 * (n_local + M5 * (2 * len(groups) - 1)) + 4.0 * n_local * M5 * n_pts
        mix_block = {
            "metric": "object mixing Msamples/s (input object-samples, whole job)",
            "value": n_total * T5 / (mix_ms * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": mix_ms,
            "ms_per_step_reduce_scatter": mix_rs_ms, "scaling": "strong",
            "config": {"workload": f"configs[4] (SURVEY cfg 5): {n_total} extent objects x {T5 / 48000:.0f} s x 48 kHz into {lay5} (M={M5}), "
                                   f"{n_local} objects/GPU in groups of 256, gains from GainCalculatorObjects at "
                                   f"{n_pts} points; HOA order 3 (16 ch) decode alongside",
                       "bus_reduce": (("node-wide bus over peer memory (earb200_bus_*): the last group's render stores the "
                                       "partial tiles into the owners' slots, deterministic sum per time slice, gather to "
                                       "every rank; bus_reduce_ms is what remains after the last render"
                                       if mbus is not None else
                                       "nccl all_reduce(sum, f32) of the [M x T] bus after the render; bus_reduce_ms is the reduce"))
                       if world > 1 else "none (1 GPU)"},
            "roofline": {"bound": "hbm", "achieved": mix_bytes / (mix_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": mix_bytes / (mix_ms * 1e-3) / 1e9 / hbm_peak, "traffic": None,
                         "peak_source": peak_src, "kernel": "render_tc_persistent_kernel (+ all_reduce when sharded)",
                         "algorithmic_bytes_per_step_per_gpu": mix_bytes},
            "bus_reduce_ms": reduce_ms,
            "bus_reduce_gbs": (4.0 * M5 * T5 / (reduce_ms * 1e-3) / 1e9) if reduce_ms else None,
            "hoa": {"design_ms": hoa_design_ms, "apply_ms": hoa_ms,
                    "apply_msamples_per_s": 16 * T5 / (hoa_ms * 1e-3) / 1e6,
                    "apply_gbs": 4.0 * T5 * (16 + M5) / (hoa_ms * 1e-3) / 1e9},
        }
        if world > 1 and not args.no_parity:
            # rank 0 renders a window of ALL 4,096 objects alone and compares with the summed bus
            Tw = 4096
            n_max = shard_range(n_total, 0, world)[1]
            xw = [torch.zeros((n_max, Tw), dtype=torch.float32, device=dev) for _ in range(world)]
            mine = torch.zeros((n_max, Tw), dtype=torch.float32, device=dev)
            mine[:n_local] = xo[:, :Tw]
            dist.all_gather(xw, mine)
            gw = [None] * world
            dist.all_gather_object(gw, gains5[:2])
            step_mix()
            torch.cuda.synchronize()
            if rank == 0:
                got = torch.empty((M5, Tw), dtype=torch.float32, device=dev)
                if mbus is not None and args.bus_gather:
                    mbus.copy_result(got.data_ptr(), Tw, 0, Tw, stream=stream)
                elif mbus is not None:
                    got.copy_(slice5[:, :Tw])
                else:
                    got.copy_(bus[:, :Tw])
                want = torch.zeros((M5, Tw), dtype=torch.float32, device=dev)
                first = True
                for r in range(world):
                    nr = gw[r].shape[1]
                    for g0, g1 in object_groups(nr, 256):
                        gr = eb.GainInterpolator(g1 - g0, M5, device=local_rank)
                        gr.set_points(samples5[:2], gw[r][:, g0:g1])
                        gr.set_accumulate(not first)
                        first = False
                        gr.process_device(0, Tw, xw[r][g0].data_ptr(), Tw, want.data_ptr(), Tw, stream=stream)
                torch.cuda.synchronize()
                err = float((got - want).abs().max() / want.abs().max())
                parity["sharded_mix_vs_one_gpu_max_over_peak"] = err
                parity["ok"] = parity["ok"] and err < 1e-6
            del xw, mine
        del xo, bus, groups, mbus

    if rank == 0:
        parity = {k: (bool(v) if isinstance(v, np.bool_) else v) for k, v in parity.items()}
        line = {
            "metric": "object gain vectors/s (9+10+3, extent)", "value": value, "unit": "vectors/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"configs[1] (SURVEY cfg 2): {n} extent objects/GPU on {LAYOUT} via the batched "
                                   "GainCalculatorObjects C-ABI (1,652-point spreading grid, depth != 0 for half)",
                       "layout": LAYOUT, "objects_per_gpu": n, "passes_per_step_per_gpu": n_passes,
                       "l2": "inputs+outputs 256 MB/step/GPU > 126 MB L2", "sharding": "objects, no collective"},
            "roofline": {"bound": "tensor", "achieved": core_tflops, "peak": bf16_peak, "unit": "TFLOP/s",
                         "frac": core_tflops / bf16_peak,
                         "traffic": ncu_traffic("extent_cell_kernel") if n >= 262144 else None,
                         "peak_source": peak_src,
                         "kernel": "extent_cell_kernel", "kernel_ms": core_ms / max(1, core_launches),
                         "algorithmic_flop_per_pass": FLOP_PER_PASS,
                         "note": "contraction flops of the reference's loop (all 1,652 points of every pass); the "
                                 "kernel culls cells of 16 points whose weights are all 0 or all 1 and evaluates "
                                 "the rest point by point -- issue-bound on that, not on the tensor pipe (DESIGN.md section 3)",
                         "weight_evals_per_s": weight_evals,
                         "kernel_share_of_step": core_ms / ms_total},
            "e2e": {"value": e2e_value, "unit": "vectors/s", "h2d_bytes_per_step": 8 * 8 * n,
                    "d2h_bytes_per_step": 2 * 4 * C_out * n, "steps": e2e_steps, "checksum": e2e_checksum,
                    "floor": e2e_floor},
            "gpu_launches": int(gpu_launches),
            "clocks": sampler.summary(),
            "parity": parity,
        }
        if objects_sustained is not None:
            line["value_sustained"] = objects_sustained["value"]
            line["sustained"] = objects_sustained
        if point_block is not None:
            line["point_sources"] = point_block
        if p5200_block is not None:
            line["extent_p5200"] = p5200_block
        if render is not None:
            line["render"] = render
        if conv_block is not None:
            line["convolver"] = conv_block
        if flow_block is not None:
            line["flow"] = flow_block
        if mix_block is not None:
            line["mix"] = mix_block
        if world == 1 and not args.no_cpu_baseline:
            from oracle import ref
            if ref.available():
                threads = host_threads()
                rate, cn, cdt = cpu_objects_throughput(objs, threads, 12.0)
                line["cpu_baseline"] = {
                    "value": rate, "unit": "vectors/s", "cores": threads, "kind": "reference",
                    "sample": f"{cn} object evaluations drawn from the {n} objects, {cdt:.1f} s wall, libear "
                              f"GainCalculatorObjects::calculate<float> (xsimd ISA {ref.extent_isa()}), "
                              "one calculator per thread"}
            else:
                line["cpu_baseline"] = {"value": None, "unit": "vectors/s", "cores": 0, "kind": "reference",
                                        "sample": "oracle/_ref/libear_ref.so not built"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0 and parity.get("ok") is False:
        raise SystemExit("bench.py: the parity check of the timed regions FAILED: " + json.dumps(parity))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--objects", type=int, default=N_OBJECTS)
    ap.add_argument("--no-render", action="store_true")
    ap.add_argument("--no-convolver", action="store_true")
    ap.add_argument("--no-mix", action="store_true")
    ap.add_argument("--no-flow", action="store_true")
    ap.add_argument("--no-points", action="store_true")
    ap.add_argument("--conv-blocks", type=int, default=56250, help="blocks per channel (cfg4: 10 min)")
    ap.add_argument("--mix-objects", type=int, default=4096, help="objects of the whole job (cfg5)")
    ap.add_argument("--mix-samples", type=int, default=2_880_000)
    ap.add_argument("--bus-gather", type=int, default=1,
                    help="1: every rank ends with the whole summed bus (all-reduce semantics, like the NCCL arm); "
                         "0: every rank ends with its time slice only (reduce-scatter semantics)")
    ap.add_argument("--exchange", default="bus", choices=["bus", "nccl"],
                    help="the exchange step of object-sharded mixing at N > 1: the node-wide bus of the C ABI (peer "
                         "stores from the render epilogue), or the comparison arm: render, then one NCCL all-reduce")
    ap.add_argument("--no-render-e2e", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiler runs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the libear check of the timed regions' output")
    ap.add_argument("--sustained", type=float, default=3.0,
                    help="seconds of back-to-back steps per block for the sustained figures (0 = off)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least three warm-up steps
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
