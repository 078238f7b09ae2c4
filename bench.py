#!/usr/bin/env python
"""bench.py -- headline benchmark of the frame-recurrent 4x EnhanceNet inference path (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload configs4|configs1]

Workload (every N, so that the N=1 line of the scaling sweep equals the single-GPU bench): BASELINE.json configs[4] -- the
throughput sweep: EnhanceNet unshaded 4x frame-recurrent, 64 independent sequences per GPU, 512x512 -> 2048x2048,
processed as 16 chunks of 4 sequences (a 64-channel 4x activation of all 64 would be 32 GiB; the chunks are bit-identical
to the unchunked result, tests/test_gpu_parity.py::test_sequence_runner_chunks_are_bit_identical).  One "step" = one
frame of the recurrence for all 64 sequences (mainVideoUnshaded.py:419-465): per chunk fused warp + space-to-depth +
concat -> 25 tensor-core convolutions -> fused skip / clamp / normalise state -> screen-space shading.  N>1: every rank
owns its own 64 sequences (sharded by sequence, no data-path collective; NCCL only for the barrier and the max-over-ranks
reduction) -> "scaling": "weak".

Prints ONE JSON line (rank 0).  value = output Mpix/s with inputs resident in HBM; e2e = the same metric through the
C-ABI pipeline with pinned HOST buffers (H2D of low + flow and D2H of the fp32 prediction of every chunk of every step
inside the timed region).  Sub-keys keep the other BASELINE configs: "configs1" (16 x 256^2, the round-1 headline),
"latency_1080p" (configs[2]), "configs3" (shaded 4 x 540x960).
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

METRIC = "4x-upscaled output Mpix/s"
UNIT = "Mpix/s"
T_FRAMES = 10
WORKLOADS = {
    "configs4": {"B": 64, "chunk": 4, "h": 512, "w": 512,
                 "name": "EnhanceNet unshaded 4x frame-recurrent throughput sweep, 512x512->2048x2048, 64 sequences per GPU "
                         "in 16 chunks of 4 (BASELINE.json configs[4]); step = 1 recurrent frame of all 64 sequences incl. "
                         "warp+s2d, 25 convs, state post-process, shading"},
    "configs1": {"B": 16, "chunk": 16, "h": 256, "w": 256,
                 "name": "EnhanceNet unshaded 4x frame-recurrent, 256x256->1024x1024, 16 sequences per GPU "
                         "(BASELINE.json configs[1]); step = 1 recurrent frame of the batch incl. warp+s2d, 25 convs, "
                         "state post-process, shading"},
}
FLOP_PER_OUT_PX_CONV64 = 2 * 64 * 64 * 9            # the dominant kernel: one 64->64 conv3x3 at 4h x 4w
FLOP_PER_OUT_PX_NET = 272232                        # algorithmic FLOPs of the whole conv stack per output px
PRECISION = os.environ.get("ISR_PRECISION", "hybrid")
DTYPE_NAMES = {"mixed": "bf16x3(trunk)+fp16(tail), fp32 accumulate", "bf16x3": "bf16x3", "bf16": "bf16", "fp16": "fp16",
               "fp16x3": "fp16x3", "mixed2": "bf16x3(trunk)+fp16x2(tail)",
               "hybrid": "fp16 (x3 split operands in the preblock + residual trunk with an 8-bit e4m3 x e5m2 correction pass, "
                         "single-pass 4x tail, hi+lo weights in the output conv), fp32 accumulate"}


def load_weights():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1350.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1350.0, 6650.0, "fallback"


def ncu_traffic(kernel, shape_key):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch from the committed `ncu --set full` summaries, keyed by
    kernel and shape (profiles/ncu_traffic.json); None when this shape has not been captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        e = json.load(open(p))[kernel][shape_key]
        return float(e["dram_bytes_read"] + e["dram_bytes_write"]), e.get("source")
    except (OSError, KeyError, ValueError, TypeError):
        return None, None


def ncu_tensor_activity():
    """Time-weighted tensor-pipe activity of one recurrent step from the committed ncu pass (profiles/r02_tensor_activity.json:
    sm__pipe_tensor_subpipe_hmma_cycles_active x gpu__time_duration per launch); a profile, not a live measurement."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_tensor_activity.json")))["configs4_chunk_4x512x512"]
        return {"conv_stack_pct": d["tensor_pipe_active_pct_time_weighted_conv_stack"],
                "whole_step_pct": d["tensor_pipe_active_pct_time_weighted_whole_step"],
                "source": "profiles/r02_tensor_activity.json (ncu, one configs[4] chunk, precision hybrid K=10)"}
    except (OSError, KeyError, ValueError):
        return None


def physical_cores():
    try:
        seen = set()
        phys = core = None
        for line in open("/proc/cpuinfo"):
            if line.startswith("physical id"):
                phys = line.split(":")[1].strip()
            elif line.startswith("core id"):
                core = line.split(":")[1].strip()
            elif not line.strip():
                if phys is not None and core is not None:
                    seen.add((phys, core))
                phys = core = None
        usable = len(os.sched_getaffinity(0))
        return max(1, min(len(seen) or usable, usable))
    except OSError:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i] == "Active" for r in self.rows)]
        busy = [v for v in sm if v > 0.5 * (max(mx) if mx else 1)] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------------
# The reference's CPU path (torch fp32 on the host cores), restated in oracle/oracle.py
def cpu_reference_leg(steps, warmup, h, w, sample_b=1, threads=None):
    """One step = one recurrent frame (warp, flatten, concat, EnhanceNet, clamp/normalise, shading) for `sample_b`
    sequences of h x w -> 4h x 4w.  Returns (Mpix/s, ms/step, threads, sample description)."""
    import torch
    from oracle import oracle as orc
    threads = threads or physical_cores()
    torch.set_num_threads(threads)
    weights = load_weights()
    low, flow, _ = orc.synthetic_gbuffer(sample_b, h, w, 4321 + 1, T=2)
    prm = orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5)
    prev = orc.postprocess(orc.initial_image(low[:, 0], 6, "input", False, 4))

    def step(prev):
        warped = orc.warp_upscale(prev, flow[:, 0], 4, special_mask=True, backend="torch")
        single = np.concatenate([low[:, 1], orc.flatten_high(warped, 4, backend="torch")], axis=1)
        recon, _ = orc.enhancenet_forward(weights, single, "bilinear", "residual", 5, backend="torch")
        state = orc.postprocess(recon, unshaded=True)
        orc.shade(state, prm)
        return state

    for _ in range(warmup):
        prev = step(prev)
    t0 = time.perf_counter()
    for _ in range(steps):
        prev = step(prev)
    dt = time.perf_counter() - t0
    mpix = sample_b * 16 * h * w * steps / dt / 1e6
    return mpix, dt / steps * 1e3, threads, (f"{sample_b} sequence(s), {steps} recurrent frames {h}x{w}->{4 * h}x{4 * w}, "
                                             f"torch {torch.__version__} fp32 CPU, {threads} threads")


def cpu_c1_leg():
    """BASELINE.json configs[0] in full: single 128x128 -> 512x512 frame, batch 1, reference PyTorch on CPU; 2 warm-up +
    5 timed frames, on all physical cores and with the reference's own torch.set_num_threads(4)
    (mainVideoUnshaded.py:174)."""
    out = {}
    for label, thr in (("physical_cores", physical_cores()), ("reference_setting_4_threads", 4)):
        mpix, ms, threads, sample = cpu_reference_leg(5, 2, 128, 128, 1, thr)
        out[label] = {"value": mpix, "unit": UNIT, "ms_per_frame": ms, "threads": threads, "sample": sample}
    return out


def run_reference(args, rank):
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    mpix, ms, cores, sample = cpu_reference_leg(args.steps, args.warmup, wl["h"], wl["w"], 1)
    line = {"impl": "reference", "metric": METRIC, "value": mpix, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["name"] + " -- CPU port, bounded sample: 1 of the %d sequences per step" % wl["B"],
                       "sequences": 1, "frames_per_step": 1, "same_config": False},
            "cpu_baseline": {"value": mpix, "unit": UNIT, "cores": cores, "logical_cpus": os.cpu_count(), "kind": "port",
                             "sample": sample, "configs0_full": cpu_c1_leg()},
            "e2e": {"value": mpix, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
def pin_to_gpu_numa_node(torch, local_rank):
    """Before any pinned allocation: bind this rank to the CPUs of the NUMA node its GPU hangs off (sysfs), so that the
    pinned staging buffers are first-touched on that node and the D2H stream does not cross the socket interconnect."""
    info = {"applied": False}
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read().strip())
        cpus = open(base + "/local_cpulist").read().strip()
        info.update({"pci": bdf, "numa_node": node, "local_cpulist": cpus})
        ids = set()
        for part in cpus.split(","):
            if "-" in part:
                a, b = part.split("-")
                ids.update(range(int(a), int(b) + 1))
            elif part:
                ids.add(int(part))
        ids &= os.sched_getaffinity(0)
        if ids:
            os.sched_setaffinity(0, ids)
            info["applied"] = True
            info["cpus_bound"] = len(ids)
    except (OSError, ValueError, AttributeError) as e:
        info["error"] = repr(e)
    return info


def run_b200(args, rank, world, local_rank):
    import ctypes
    from argparse import Namespace

    import torch
    import torch.distributed as dist

    import isosurfacesuperresolution_b200 as isr
    from isosurfacesuperresolution_b200 import _lib, ops, sharding, synthetic   # the product leg never imports oracle/

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    affinity = pin_to_gpu_numa_node(torch, local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    weights = load_weights()
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6,
                            Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    net = net.to(dev).eval()
    net.precision = PRECISION
    shd = isr.ScreenSpaceShading(dev)
    shd.fov(30)
    shd.ambient_light_color(np.array([0.1, 0.1, 0.1])); shd.diffuse_light_color(np.array([1.0, 1.0, 1.0]))
    shd.specular_light_color(np.array([0.2, 0.2, 0.2])); shd.specular_exponent(16)
    shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0]))
    shd.ambient_occlusion(0.5); shd.background(np.array([0.0, 0.0, 0.0]))
    handle = net._native()
    T = T_FRAMES

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x, dtype=torch.float64):
        t = torch.tensor([x], device=dev, dtype=dtype)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- one workload: device-resident value + end-to-end through the host pipeline --------------------------------
    def measure(wl, steps, warmup, e2e_variants, sample_clocks):
        B, chunk, h, w = wl["B"], wl["chunk"], wl["h"], wl["w"]
        if os.environ.get("ISR_BENCH_CHUNK"):      # experiment: another chunking of the same workload
            chunk = int(os.environ["ISR_BENCH_CHUNK"])
        n_chunks = B // chunk
        H, W = 4 * h, 4 * w
        out_px_per_step = B * H * W
        # 2 distinct sequences generated on the host and tiled to one chunk (generation cost only), per-rank seed; every
        # chunk of the rank reads the same low / flow tensors and owns its own recurrent state
        low_np, flow_np, _ = synthetic.synthetic_gbuffer(2, h, w, sharding.rank_seed(4321 + 1, rank), T=T)
        low = torch.from_numpy(np.tile(low_np, ((chunk + 1) // 2, 1, 1, 1, 1))[:chunk]).to(dev)
        flow = torch.from_numpy(np.tile(flow_np, ((chunk + 1) // 2, 1, 1, 1, 1))[:chunk]).to(dev)
        res = {}
        with torch.no_grad():
            inp0 = low[:, 0].contiguous()
            states = []
            for c in range(n_chunks):      # frame 0 (initial image), untimed; leaves each chunk's state on the device
                prev = isr.initialImage(inp0, 6, "input", False, 4)
                _, _, st = net(inp0, prev, None, return_state=True)
                states.append(st)
            lows = [low[:, j].contiguous() for j in range(T)]
            flows = [flow[:, j].contiguous() for j in range(T)]

            def step(j):
                fj = 1 + (j % (T - 1))
                for c in range(n_chunks):
                    # the state is updated in place: the packing kernel reads it at the start of the frame, the output
                    # conv's epilogue writes the new one at the end (same stream)
                    recon, _, st = handle.forward_recurrent(lows[fj], states[c], flows[fj - 1], True, PRECISION,
                                                            want_raw=False, want_state=True, out_state=states[c])
                    rgb = shd(st)
                return recon, rgb

            for j in range(warmup):
                step(j)
            sampler = ClockSampler(local_rank)
            sync_all()
            if rank == 0 and sample_clocks:
                sampler.start()
            n0 = _lib.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for j in range(steps):
                step(j)
            e1.record()
            torch.cuda.synchronize()
            res["launches"] = _lib.launch_count() - n0
            ms_total = max_over_ranks(e0.elapsed_time(e1))
            if world > 1:
                dist.barrier()
            res["clocks"] = sampler.stop() if (rank == 0 and sample_clocks) else None
            res["ms_per_step"] = ms_total / steps
            res["value"] = world * out_px_per_step * steps / (ms_total * 1e-3) / 1e6

            # ---- e2e: the pipelined host-buffer API (isr_pipe_*): pinned HOST inputs and outputs, H2D of low + flow and
            # D2H of the result for EVERY chunk of every step inside the timed region; the copies of neighbouring chunks
            # overlap the kernels (three streams, depth-2 staging); one pipeline serves all chunks (isr_pipe_set_state)
            nbuf = min(steps, T - 1)
            low_h = [lows[1 + (j % (T - 1))].cpu().pin_memory() for j in range(nbuf)]
            flow_h = [flows[j % (T - 1)].cpu().pin_memory() for j in range(nbuf)]
            low0_h = lows[0].cpu().pin_memory()
            pipe = ops.RecurrentPipeline(handle, 5, 6, True, "input", chunk, h, w, PRECISION, depth=2)
            e2e_steps = max(3, steps)
            h2d = B * 7 * h * w * 4
            e2e = {}
            for variant in e2e_variants:
                if variant == "fp32":
                    bufs = [torch.empty((chunk, 6, H, W), dtype=torch.float32).pin_memory() for _ in range(3)]
                    submit = pipe.submit
                    d2h = B * 6 * H * W * 4
                elif variant == "fp16_out":
                    pipe.enable_f16()
                    bufs = [torch.empty((chunk, 6, H, W), dtype=torch.float16).pin_memory() for _ in range(3)]
                    submit = pipe.submit_f16
                    d2h = B * 6 * H * W * 2
                else:
                    pipe.enable_display(shd.params22())
                    bufs = [torch.empty((chunk, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(3)]
                    submit = pipe.submit_display
                    d2h = B * H * W * 3
                k = 0
                for c in range(n_chunks):                      # frame 0 of every chunk (initial image), untimed
                    pipe.set_state(states[c])
                    submit(low0_h, None, bufs[k % 3]); k += 1
                for c in range(n_chunks):
                    pipe.set_state(states[c])
                    submit(low_h[0], flow_h[0], bufs[k % 3]); k += 1
                pipe.wait()
                sync_all()
                t0 = time.perf_counter()
                for j in range(e2e_steps):
                    for c in range(n_chunks):
                        pipe.set_state(states[c])
                        submit(low_h[j % nbuf], flow_h[j % nbuf], bufs[k % 3]); k += 1
                pipe.wait()
                dt = max_over_ranks(time.perf_counter() - t0)
                last = bufs[(k - 1) % 3]
                e2e[variant] = {"value": world * out_px_per_step * e2e_steps / dt / 1e6, "unit": UNIT,
                                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                                "d2h_gbs_all_ranks": world * d2h * e2e_steps / dt / 1e9,
                                "output_checksum": float(last.double().abs().mean())}
                del bufs
            pipe.close()
            res["e2e"] = e2e
            res["handles"] = (B, chunk, h, w, H, W, out_px_per_step, low_h, flow_h, states)
        return res

    wl = WORKLOADS[args.workload]
    main_res = measure(wl, args.steps, args.warmup, ("fp32", "fp16_out", "display"), True)
    B, chunk, h, w, H, W, out_px_per_step, low_h, flow_h, states = main_res.pop("handles")

    with torch.no_grad():
        # ---- the same step through the synchronous call (copies and kernels serialised), one chunk, for reference ---
        staging = torch.empty(chunk * 7 * h * w + chunk * 6 * H * W, device=dev, dtype=torch.float32)
        recon_one = torch.empty((chunk, 6, H, W), dtype=torch.float32).pin_memory()
        _, ws_ptr, ws_bytes = handle.workspace(chunk, h, w, PRECISION)
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

        def sync_step(j):
            k = j % len(low_h)
            _lib.check(_lib.lib().isr_net_forward_recurrent_host(
                handle._net, ctypes.c_void_p(low_h[k].data_ptr()), ctypes.c_void_p(flow_h[k].data_ptr()),
                ctypes.c_void_p(states[0].data_ptr()), 1, 2, 5, 6, 1, ctypes.c_void_p(recon_one.data_ptr()), chunk, h, w,
                _lib.PREC[PRECISION], ctypes.c_void_p(ws_ptr), ws_bytes, ctypes.c_void_p(staging.data_ptr()), stream))

        sync_step(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for j in range(3):
            sync_step(j)
        torch.cuda.synchronize()
        e2e_sync_value = chunk * H * W * 3 / (time.perf_counter() - t0) / 1e6
        del staging

        # ---- the box's device-to-host ceiling: every rank streams 403 MB blocks into pinned memory AT THE SAME TIME, into
        # three rotating buffers like the pipeline does (a single reused buffer flatters the host side: 120 instead of
        # ~100 GB/s aggregate at 8 ranks)
        n_probe = recon_one.numel()
        src = torch.empty(n_probe, device=dev)
        dsts = [recon_one.view(-1)] + [torch.empty(n_probe).pin_memory() for _ in range(2)]
        for d in dsts:
            d.copy_(src, non_blocking=True)
        sync_all()
        t0 = time.perf_counter()
        n_copies = 16
        for k in range(n_copies):
            dsts[k % 3].copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        my_s = time.perf_counter() - t0
        # the concurrent aggregate: all ranks' bytes over the time of the SLOWEST rank (they start at the same barrier).
        # The sum of the ranks' own rates is reported too: it overstates the box when the link is shared unevenly (ranks
        # that finish early leave their share to the others)
        d2h_ceiling = world * n_copies * n_probe * 4 / max_over_ranks(my_s) / 1e9
        d2h_ceiling_sum_of_rates = sum_over_ranks(n_copies * n_probe * 4 / my_s / 1e9)
        del src, dsts

        # ---- roofline of the dominant kernel (64->64 conv at 4h x 4w), CUDA events on the launching stream ---------
        conv_ms = handle.time_conv(chunk, h, w, PRECISION, scale=4, iters=10)
        peak_tf, peak_tf_sustained, peak_hbm, which = peaks()
        conv_px = chunk * H * W
        flops = FLOP_PER_OUT_PX_CONV64 * conv_px
        achieved = flops / (conv_ms * 1e-3) / 1e12
        traffic, traffic_src = ncu_traffic("conv3x3_pair_kernel<0,64,0,1>", f"{chunk}x{h}x{w}")
        stack_tflops = FLOP_PER_OUT_PX_NET * out_px_per_step / (main_res["ms_per_step"] * 1e-3) / 1e12

    del states
    torch.cuda.empty_cache()

    # ---- sub-keys: the other BASELINE configs ----------------------------------------------------------------------
    extra = {}
    if world == 1 and not args.no_extras and args.workload != "configs1":
        r1 = measure(WORKLOADS["configs1"], min(args.steps, 20), 3, ("fp32",), False)
        r1.pop("handles")
        extra["configs1"] = {"workload": WORKLOADS["configs1"]["name"], "value": r1["value"], "unit": UNIT,
                             "ms_per_step": r1["ms_per_step"], "e2e": r1["e2e"]["fp32"], "gpu_launches": int(r1["launches"])}
        torch.cuda.empty_cache()
        extra["configs3"] = bench_configs3(isr, torch, dev)
        torch.cuda.empty_cache()
    latency = None
    if world == 1 and not args.no_latency:
        latency = bench_latency(isr, ops, synthetic, torch, dev, handle, shd)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    e2e = main_res["e2e"]
    line = {
        "metric": METRIC, "value": main_res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": main_res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": DTYPE_NAMES.get(PRECISION, PRECISION), "data": "synthetic",
        "config": {"workload": wl["name"], "sequences_per_gpu": B, "chunk": chunk, "frames": T_FRAMES, "low_res": [h, w],
                   "precision": PRECISION,
                   "l2": "inputs_larger_than_L2 (per-chunk activations ~7 GB >> 126 MB L2)", "sharding": "by sequence"},
        "e2e": dict(e2e["fp32"], api="isr_pipe_set_state / isr_pipe_submit / isr_pipe_wait, depth 2 (pinned host buffers; "
                                     "copies of neighbouring chunks overlap the kernels)",
                    serialised_value=e2e_sync_value, serialised_api="isr_net_forward_recurrent_host (one chunk)",
                    d2h_ceiling_gbs=d2h_ceiling,
                    d2h_ceiling_how="all ranks copy 16 blocks of 403 MB device->pinned host from one barrier, three rotating buffers "
                                    "per rank: all ranks' bytes / the slowest rank's time",
                    d2h_ceiling_sum_of_rank_rates_gbs=d2h_ceiling_sum_of_rates,
                    d2h_frac_of_ceiling=e2e["fp32"]["d2h_gbs_all_ranks"] / d2h_ceiling if d2h_ceiling else None,
                    affinity=affinity,
                    fp16_out=dict(e2e["fp16_out"], api="isr_pipe_submit_f16: the host result is the full prediction rounded "
                                                       "to fp16 (12 instead of 24 B per output pixel)"),
                    display=dict(e2e["display"], api="isr_pipe_submit_display: the host result is the shaded 8-bit RGB "
                                                     "frame (B x H x W x 3)")),
        "gpu_launches": int(main_res["launches"]),
        "clocks": main_res["clocks"],
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved / peak_tf, "traffic": traffic, "traffic_source": traffic_src,
                     "traffic_unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full)",
                     "algorithmic_bytes": 2 * 128 * conv_px, "peak_source": which + " bf16_tflops (burst)",
                     "kernel": "conv3x3_pair_kernel<0,64,0,1> (tcgen05 cta_group::2) 64->64 @%dx%d x%d, algorithmic 73,728 "
                               "FLOP/px" % (H, W, chunk), "ms_per_launch": conv_ms,
                     "whole_conv_stack": {"algorithmic_tflops": stack_tflops, "frac_of_burst": stack_tflops / peak_tf,
                                          "frac_of_sustained": stack_tflops / peak_tf_sustained,
                                          "how": "272,232 algorithmic FLOP per output px x px per step / ms_per_step "
                                                 "(the step also holds packing and shading)",
                                          "tensor_pipe_active": ncu_tensor_activity()}},
    }
    line.update(extra)
    if latency is not None:
        line["latency_1080p"] = latency
    if world == 1 and not args.no_cpu_baseline:
        mpix, cms, cores, sample = cpu_reference_leg(3, 1, h, w, 1)
        line["cpu_baseline"] = {"value": mpix, "unit": UNIT, "cores": cores, "logical_cpus": os.cpu_count(),
                                "kind": "port", "sample": sample, "ms_per_step": cms, "configs0_full": cpu_c1_leg()}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_configs3(isr, torch, dev):
    """BASELINE.json configs[3]: shaded (RGB, mainVideo.py variant) 56 -> 3, 4 x 540x960 -> 2160x3840, one recurrent step
    (3-channel previous frame, special_mask=False, clamp state), device-resident."""
    from argparse import Namespace
    torch.manual_seed(1234)
    net = isr.createNetwork("EnhanceNet", 4, 56, [0, 1, 2], 3,
                            Namespace(upsample="bilinear", reconType="residual", useBN=False)).to(dev).eval()
    net.precision = PRECISION
    B, h, w = 4, 540, 960
    g = torch.Generator(device=dev).manual_seed(5)
    low = torch.rand(B, 8, h, w, device=dev, generator=g)
    flow = (torch.rand(B, 2, h, w, device=dev, generator=g) - 0.5) * 0.04
    state = torch.rand(B, 3, 4 * h, 4 * w, device=dev, generator=g)
    with torch.no_grad():
        for _ in range(3):
            pred, _, state = net(low, state, flow, special_mask=False, return_state=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            pred, _, state = net(low, state, flow, special_mask=False, return_state=True)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    return {"workload": "EnhanceNet shaded (RGB) 4x, 960x540->3840x2160, batch 4 (BASELINE.json configs[3]), one recurrent step",
            "value": B * 16 * h * w / ms / 1e3, "unit": UNIT, "ms_per_step": ms, "steps": 10,
            "finite": bool(torch.isfinite(pred).all())}


def bench_latency(isr, ops, synthetic, torch, dev, handle, shd):
    """configs[2]: interactive 1080p frame latency (1 x 270x480 -> 1080x1920), one CUDA graph per frame: fused warp +
    space-to-depth + concat, the conv stack, fused state, composition incl. shading."""
    lh, lw = 270, 480
    l_low, l_flow, _ = synthetic.synthetic_gbuffer(1, lh, lw, 4321 + 2, T=2)
    l_low_t, l_flow_t = torch.from_numpy(l_low).to(dev), torch.from_numpy(l_flow).to(dev)
    render = torch.rand(1, 12, lh, lw, device=dev)
    render[:, 3:8] = l_low_t[:, 1]
    render[:, 8:10] = l_flow_t[:, 0]
    l_inp = l_low_t[:, 1].contiguous()
    l_flow0 = l_flow_t[:, 0].contiguous()
    l_state = isr.initialImage(l_low_t[:, 0].contiguous(), 6, "input", False, 4)
    params = shd.params22()

    def timed_graph(fn):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            display = fn()
        for _ in range(20):
            graph.replay()
        torch.cuda.synchronize()
        times = []
        for _ in range(200):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            graph.replay()
            a1.record()
            a1.synchronize()
            times.append(a0.elapsed_time(a1))
        times.sort()
        return times, display

    with torch.no_grad():
        def frame():
            _, _, st = handle.forward_recurrent(l_inp, l_state, l_flow0, True, PRECISION, want_raw=False,
                                                want_state=True, out_state=l_state)
            return ops.compose(render, st, params, True, False, 0.0)

        times, display = timed_graph(frame)

        # the same frame starting from the renderer's interleaved H x W x 12 buffer: ingest kernel
        # (slicing, mask remap, hole mask, GUI frame) + device flow in-painting replace the reference's
        # GPU -> CPU cv.inpaint -> GPU round trip (loadedmodel.py:77-82)
        wire = torch.rand(1, lh, lw, 12, device=dev)
        wire[..., 3] = (l_low_t[0, 1, 0] > 0).float()
        wire[..., 4:8] = l_low_t[0, 1, 1:5].permute(1, 2, 0)
        wire[..., 8:10] = (l_flow_t[0, 0] * (l_low_t[0, 1, 0:1] > 0).float()).permute(1, 2, 0)
        wire = wire.contiguous()
        l_state2 = l_state.clone()

        def frame_from_render():
            inp, fl, hole, img = ops.ingest_gbuffer(wire, "unshaded", interleaved=True, want_image12=True)
            fl = ops.inpaint_flow(fl, hole)
            _, _, st = handle.forward_recurrent(inp, l_state2, fl, True, PRECISION, want_raw=False, want_state=True,
                                                out_state=l_state2)
            return ops.compose(img, st, params, True, False, 0.0)

        times2, display2 = timed_graph(frame_from_render)
    from_render = {"p50_ms": times2[len(times2) // 2], "p99_ms": times2[int(len(times2) * 0.99) - 1],
                   "adds": "isr_ingest_gbuffer (interleaved 270x480x12 -> input, flow, hole mask, GUI frame) + "
                           "isr_inpaint_flow (push-pull, 19 small launches) in the same graph",
                   "display_checksum": float(display2.double().abs().mean())}
    return {"p50_ms": times[len(times) // 2], "p99_ms": times[int(len(times) * 0.99) - 1], "frames": len(times),
            "config": "BASELINE.json configs[2]: 1 x 480x270 -> 1920x1080, per frame: warp+s2d+concat, EnhanceNet, "
                      "state post-process, 12-channel composition incl. Phong+AO shading; inputs on the device",
            "launch": "one CUDA graph replay per frame", "output_Mpix_per_s_at_p50":
                1080 * 1920 / (times[len(times) // 2] * 1e-3) / 1e6,
            "display_checksum": float(display.double().abs().mean()), "from_render": from_render}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("ISR_BENCH_WORKLOAD", "configs4"), choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the configs1 / configs3 sub-keys")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
