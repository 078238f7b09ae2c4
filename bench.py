#!/usr/bin/env python
"""bench.py -- headline benchmark of the frame-recurrent 4x EnhanceNet inference path (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (N=1): BASELINE.json configs[1] -- EnhanceNet unshaded 4x frame-recurrent, synthetic 10-frame sequences
256x256 -> 1024x1024, batch 16 sequences per GPU.  One "step" = one frame of the recurrence for the whole batch:
fused warp + space-to-depth + concat -> 25 tensor-core convolutions (+2 upsamples) -> fused skip / clamp / normalise state
-> screen-space shading.  N>1: every rank owns its own 16 sequences (sharded by sequence, no data-path collective;
NCCL only for the barrier and the max-over-ranks reduction) -> "scaling": "weak".

Prints ONE JSON line (rank 0).  value = output Mpix/s with inputs resident in HBM; e2e = same metric through the C-ABI
call with pinned HOST buffers (H2D of low+flow and D2H of the prediction inside the timed region).
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
B_SEQ, T_FRAMES, H_LOW, W_LOW = 16, 10, 256, 256
FLOP_PER_OUT_PX_CONV64 = 2 * 64 * 64 * 9            # the dominant kernel: one 64->64 conv3x3 at 4h x 4w
PRECISION = os.environ.get("ISR_PRECISION", "hybrid")


def load_weights():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 6650.0, "fallback"


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
def cpu_reference_leg(steps, warmup, sample_b=1):
    """The reference's CPU path (torch fp32 on the host cores) restated in oracle/oracle.py: one step = one recurrent
    frame (warp, flatten, concat, EnhanceNet, clamp/normalise, shading) for `sample_b` sequences of the same 256x256
    workload.  Returns (Mpix/s, ms/step, cores, sample description)."""
    import torch
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    weights = load_weights()
    low, flow, _ = orc.synthetic_gbuffer(sample_b, H_LOW, W_LOW, 4321 + 1, T=2)
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
    mpix = sample_b * 16 * H_LOW * W_LOW * steps / dt / 1e6
    return mpix, dt / steps * 1e3, cores, (f"{sample_b} of {B_SEQ} sequences, {steps} recurrent frames {H_LOW}x{W_LOW}->"
                                          f"{4 * H_LOW}x{4 * W_LOW}, torch {torch.__version__} fp32 CPU")


def run_reference(args, rank):
    if rank != 0:
        return
    mpix, ms, cores, sample = cpu_reference_leg(args.steps, args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": mpix, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "EnhanceNet unshaded 4x frame-recurrent 256x256->1024x1024 (configs[1]), CPU sample",
                       "sequences": 1, "frames_per_step": 1},
            "cpu_baseline": {"value": mpix, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": mpix, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, local_rank):
    import ctypes
    from argparse import Namespace

    import torch
    import torch.distributed as dist

    import isosurfacesuperresolution_b200 as isr
    from isosurfacesuperresolution_b200 import _lib, ops, sharding, synthetic   # the product leg never imports oracle/

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
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

    B, T, h, w = B_SEQ, T_FRAMES, H_LOW, W_LOW
    # 2 distinct sequences generated on the host, tiled to B (generation cost only), per-rank seed
    low_np, flow_np, _ = synthetic.synthetic_gbuffer(2, h, w, sharding.rank_seed(4321 + 1, rank), T=T)
    low = torch.from_numpy(np.tile(low_np, (B // 2, 1, 1, 1, 1))).to(dev)
    flow = torch.from_numpy(np.tile(flow_np, (B // 2, 1, 1, 1, 1))).to(dev)
    H, W = 4 * h, 4 * w
    out_px_per_step = B * H * W

    with torch.no_grad():
        # frame 0 (initial image), untimed; leaves the recurrent state on the device
        inp0 = low[:, 0].contiguous()
        prev = isr.initialImage(inp0, 6, "input", False, 4)
        _, _, state = net(inp0, prev, None, return_state=True)
        handle = net._native()
        state2 = torch.empty_like(state)

        def step(j):
            nonlocal state, state2
            fj = 1 + (j % (T - 1))
            recon, _, st = handle.forward_recurrent(low[:, fj], state, flow[:, fj - 1], True, PRECISION,
                                                    want_raw=False, want_state=True, out_state=state2)
            rgb = shd(st)
            state, state2 = st, state
            return recon, rgb

        def sync_all():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()

        for j in range(args.warmup):
            step(j)
        sampler = ClockSampler(local_rank)
        sync_all()
        if rank == 0:
            sampler.start()
        n0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for j in range(args.steps):
            step(j)
        e1.record()
        torch.cuda.synchronize()
        launches = _lib.launch_count() - n0
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        clocks = sampler.stop() if rank == 0 else None
        ms_total = float(ms.item())
        value = world * out_px_per_step * args.steps / (ms_total * 1e-3) / 1e6

        # ---- e2e: the pipelined host-buffer API (isr_pipe_*): pinned HOST inputs and outputs, H2D of low + flow and D2H
        # of the fp32 prediction for EVERY step inside the timed region; copies of neighbouring frames overlap the kernels
        nbuf = min(args.steps, T - 1)
        low_h = [low[:, 1 + (j % (T - 1))].contiguous().cpu().pin_memory() for j in range(nbuf)]
        flow_h = [flow[:, (j % (T - 1))].contiguous().cpu().pin_memory() for j in range(nbuf)]
        low0_h = low[:, 0].contiguous().cpu().pin_memory()
        recon_h = [torch.empty((B, 6, H, W), dtype=torch.float32).pin_memory() for _ in range(3)]
        pipe = ops.RecurrentPipeline(handle, 5, 6, True, "input", B, h, w, PRECISION, depth=2)
        pipe.submit(low0_h, None, recon_h[0])                      # frame 0 (initial image), untimed
        for j in range(2):
            pipe.submit(low_h[j % nbuf], flow_h[j % nbuf], recon_h[(j + 1) % 3])
        pipe.wait()
        e2e_steps = max(3, args.steps)
        sync_all()
        t0 = time.perf_counter()
        for j in range(e2e_steps):
            pipe.submit(low_h[j % nbuf], flow_h[j % nbuf], recon_h[j % 3])
        pipe.wait()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e_value = world * out_px_per_step * e2e_steps / float(dt.item()) / 1e6
        e2e_checksum = float(recon_h[(e2e_steps - 1) % 3].double().abs().mean())
        # the same pipeline in its display form: the host receives the shaded 8-bit RGB frame (what the reference's GUI /
        # video writers consume, mainGUI.py:507) instead of the fp32 prediction: 3 instead of 24 bytes per output pixel
        pipe.enable_display(shd.params22())
        rgb_h = [torch.empty((B, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(3)]
        for j in range(2):
            pipe.submit_display(low_h[j % nbuf], flow_h[j % nbuf], rgb_h[j % 3])
        pipe.wait()
        sync_all()
        t0 = time.perf_counter()
        for j in range(e2e_steps):
            pipe.submit_display(low_h[j % nbuf], flow_h[j % nbuf], rgb_h[j % 3])
        pipe.wait()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e_display_value = world * out_px_per_step * e2e_steps / float(dt.item()) / 1e6
        e2e_display_checksum = float(rgb_h[(e2e_steps - 1) % 3].double().mean())
        pipe.close()
        h2d = B * 7 * h * w * 4
        d2h = B * 6 * H * W * 4

        # the same step through the synchronous call (copies and kernels serialised), for reference
        staging = torch.empty(B * 7 * h * w + B * 6 * H * W, device=dev, dtype=torch.float32)
        _, ws_ptr, ws_bytes = handle.workspace(B, h, w, PRECISION)
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

        def sync_step(j):
            k = j % nbuf
            _lib.check(_lib.lib().isr_net_forward_recurrent_host(
                handle._net, ctypes.c_void_p(low_h[k].data_ptr()), ctypes.c_void_p(flow_h[k].data_ptr()),
                ctypes.c_void_p(state.data_ptr()), 1, 2, 5, 6, 1, ctypes.c_void_p(recon_h[0].data_ptr()), B, h, w,
                _lib.PREC[PRECISION], ctypes.c_void_p(ws_ptr), ws_bytes, ctypes.c_void_p(staging.data_ptr()), stream))

        sync_step(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for j in range(3):
            sync_step(j)
        torch.cuda.synchronize()
        e2e_sync_value = out_px_per_step * 3 / (time.perf_counter() - t0) / 1e6

        # ---- roofline of the dominant kernel (64->64 conv at 4h x 4w), CUDA events on the launching stream ---------
        conv_ms = handle.time_conv(B, h, w, PRECISION, scale=4, iters=10)
        peak_tf, peak_hbm, which = peaks()
        flops = FLOP_PER_OUT_PX_CONV64 * out_px_per_step
        achieved = flops / (conv_ms * 1e-3) / 1e12

        # ---- configs[2]: interactive 1080p frame latency (1 x 270x480 -> 1080x1920), one CUDA graph per frame:
        # fused warp + space-to-depth + concat, 26 conv launches, 2 upsamples, fused state, composition incl. shading
        latency = None
        if world == 1 and not args.no_latency:
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

            def frame():
                _, _, st = handle.forward_recurrent(l_inp, l_state, l_flow0, True, PRECISION, want_raw=False,
                                                    want_state=True, out_state=l_state)
                return ops.compose(render, st, params, True, False, 0.0)

            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    frame()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                display = frame()
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

            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    frame_from_render()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph2):
                display2 = frame_from_render()
            for _ in range(20):
                graph2.replay()
            torch.cuda.synchronize()
            times2 = []
            for _ in range(200):
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                graph2.replay()
                a1.record()
                a1.synchronize()
                times2.append(a0.elapsed_time(a1))
            times2.sort()
            from_render = {"p50_ms": times2[len(times2) // 2], "p99_ms": times2[int(len(times2) * 0.99) - 1],
                           "adds": "isr_ingest_gbuffer (interleaved 270x480x12 -> input, flow, hole mask, GUI frame) + "
                                   "isr_inpaint_flow (push-pull, 19 small launches) in the same graph",
                           "display_checksum": float(display2.double().abs().mean())}
            latency = {"p50_ms": times[len(times) // 2], "p99_ms": times[int(len(times) * 0.99) - 1], "frames": len(times),
                       "config": "BASELINE.json configs[2]: 1 x 480x270 -> 1920x1080, per frame: warp+s2d+concat, EnhanceNet, "
                                 "state post-process, 12-channel composition incl. Phong+AO shading; inputs on the device",
                       "launch": "one CUDA graph replay per frame", "output_Mpix_per_s_at_p50":
                           1080 * 1920 / (times[len(times) // 2] * 1e-3) / 1e6,
                       "display_checksum": float(display.double().abs().mean()), "from_render": from_render}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"mixed": "bf16x3(trunk)+fp16(tail), fp32 accumulate", "bf16x3": "bf16x3", "bf16": "bf16",
                  "fp16": "fp16", "fp16x3": "fp16x3", "mixed2": "bf16x3(trunk)+fp16x2(tail)",
                  "hybrid": "fp16 (x3 split in preblock + 4 residual blocks, split weights in the other 6, single-pass tail), "
                            "fp32 accumulate"}.get(PRECISION, PRECISION),
        "data": "synthetic",
        "config": {"workload": "EnhanceNet unshaded 4x frame-recurrent, 256x256->1024x1024, 16 sequences per GPU "
                               "(BASELINE.json configs[1]); step = 1 recurrent frame of the batch incl. warp+s2d, "
                               "25 convs, state post-process, shading",
                   "sequences_per_gpu": B, "frames": T, "low_res": [h, w], "precision": PRECISION,
                   "l2": "inputs_larger_than_L2 (per-step activations ~7 GB >> 126 MB L2)", "sharding": "by sequence"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "api": "isr_pipe_submit / isr_pipe_wait, depth 2 (pinned host buffers; copies of "
                                             "neighbouring frames overlap the kernels)",
                "serialised_value": e2e_sync_value, "serialised_api": "isr_net_forward_recurrent_host",
                "output_checksum": e2e_checksum,
                "display": {"value": e2e_display_value, "unit": UNIT, "d2h_bytes_per_step": B * H * W * 3,
                            "api": "isr_pipe_submit_display: the host result is the shaded 8-bit RGB frame (B x H x W x 3)",
                            "output_checksum": e2e_display_checksum}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved / peak_tf,
                     # dram__bytes_read.sum + dram__bytes_write.sum of this launch, one `ncu --set full` capture
                     # (profiles/r01_ncu_conv_full_v8_two_issuers.json: 2.233 GB + 2.101 GB); algorithmic: 2 x 2.147 GB
                     "traffic": 4.334e9 if (B, h, w) == (16, 256, 256) else None,
                     "traffic_unit": "bytes per launch (ncu)", "algorithmic_bytes": 2 * 128 * out_px_per_step,
                     "peak_source": which + " bf16_tflops (burst)",
                     "kernel": "conv3x3_pair_kernel<0> (tcgen05 cta_group::2) 64->64 @1024x1024 x16, "
                               "algorithmic 73,728 FLOP/px", "ms_per_launch": conv_ms},
    }
    if latency is not None:
        line["latency_1080p"] = latency
    if world == 1 and not args.no_cpu_baseline:
        mpix, cms, cores, sample = cpu_reference_leg(steps=3, warmup=1)
        line["cpu_baseline"] = {"value": mpix, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                                "ms_per_step": cms}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
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
