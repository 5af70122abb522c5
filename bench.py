#!/usr/bin/env python3
"""bench.py — primary-ray throughput of the ATRay render path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--fp strict|fast]

Workload (config.workload): BASELINE.json configs[3] — the 139 264-face Dragon stand-in (Assets/Dragon.obj is missing
from the reference checkout; tools/gen_standin_obj.py, spec "standin-v1"), 3840×2160, 1 spp, bounce 1, default camera
(app.cpp:88), tree SAH / 300 (app.cpp:76-77). A "step" = one complete frame: ray generation → traversal →
intersection → {t, prim, BGRX} stores. N > 1: the frame is split into interleaved scanline bands, one rank per GPU,
every rank storing its bands straight into GPU 0's frame over NVLink (strong scaling, no collective).

value   = W·H·spp·K / (sum of the K per-step CUDA-event durations, max over ranks)   [frame resident in HBM]
e2e     = same frames through the reference-facing call (start_render_from_camera / wait_… → atray_render_begin/poll)
          with the host bitmap as output: kernel + device→host copy inside the timed region
--impl reference: the reference's own CPU renderer (oracle/_ref, start_render_from_camera with all host threads)."""
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

W, H, SPP, AA = 3840, 2160, 1, 0
BAND_ROWS = 8
METRIC = "Mrays/s primary rays (Dragon 139k tris, 4K)"
WORKLOAD = "C4: Dragon stand-in (standin-v1, 139264 faces) 3840x2160 1spp primary rays, SAH cap 300"
MESH_CACHE = os.environ.get("ATRAY_MESH_CACHE", "/tmp/atray_meshes")
MESH, RESCALE, CONFIG = "standin", 0.0, "c4"
ASSETS = os.path.join(ROOT, "oracle", "_ref", "assets")     # the reference's own .obj files, staged by oracle/ref/build_ref.sh
C5_SPP, C5_CHUNK = 64, 4096

# BASELINE.json configs[0..4]. The headline (and the default) is c4; c1-c3 are the smaller frames, c5 the divergence stress.
CONFIGS = {
    "c1": dict(mesh="Simple", rescale=40.0, w=640, h=360, spp=1, aa=0,
               workload="C1: Simple.obj (0 faces: every ray misses) 640x360 1spp primary rays"),
    "c2": dict(mesh="Monkey", rescale=40.0, w=1920, h=1080, spp=1, aa=0,
               workload="C2: Monkey.obj (3936 faces) 1920x1080 1spp primary rays, SAH cap 300"),
    "c3": dict(mesh="Deer", rescale=40.0, w=1920, h=1080, spp=16, aa=1,
               workload="C3: Deer.obj (1570 faces) 1920x1080 16spp jittered primary rays, SAH cap 300"),
    "c4": dict(mesh="standin", rescale=0.0, w=3840, h=2160, spp=1, aa=0, workload=WORKLOAD),
    "c5": dict(mesh="standin", rescale=0.0, w=3840, h=2160, spp=C5_SPP, aa=0,
               workload="C5: Dragon stand-in (standin-v1, 139264 faces) 3840x2160 x 64 incoherent random-direction rays per pixel "
                        "(device-generated per-ray PCG32 stream, SURVEY 8d), SAH cap 300"),
}


def apply_config(name):
    global W, H, SPP, AA, WORKLOAD, MESH, RESCALE, CONFIG, METRIC
    c = CONFIGS[name]
    W, H, SPP, AA, WORKLOAD, MESH, RESCALE, CONFIG = c["w"], c["h"], c["spp"], c["aa"], c["workload"], c["mesh"], c["rescale"], name
    if name != "c4":
        METRIC = "Mrays/s (%s)" % name.upper()


def mesh_path():
    if MESH == "standin":
        from tools import gen_standin_obj
        return gen_standin_obj.ensure(os.path.join(MESH_CACHE, "standin_272x256.obj"))
    p = os.path.join(ASSETS, MESH + ".obj")
    if not os.path.exists(p):
        raise SystemExit("config %s needs %s (staged by oracle/ref/build_ref.sh where /root/reference exists)" % (CONFIG, p))
    return p


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "MEASURED_PEAKS.json (burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and clock-event (throttle) reasons DURING the timed region (B200_PROFILING.md). A frame takes a third of a
    millisecond, so the sampler reads NVML directly (≈1 kHz); `nvidia-smi` (≈50 ms per query) is the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, gpu_index):
        self.idx, self.sm, self.mx, self.reasons, self.stop_flag, self.th, self.source = gpu_index, [], [], set(), False, None, None
        self.nv, self.h = None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(self._nvml_index(pynvml, gpu_index))
            self.source = "nvml"
        except Exception:
            self.nv, self.source = None, "nvidia-smi"

    @staticmethod
    def _nvml_index(nv, cuda_index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            tok = vis.split(",")[cuda_index].strip()
            if tok.isdigit():
                return int(tok)
            for i in range(nv.nvmlDeviceGetCount()):
                u = nv.nvmlDeviceGetUUID(nv.nvmlDeviceGetHandleByIndex(i))
                u = u.decode() if isinstance(u, bytes) else u
                if u.startswith(tok) or tok in u:
                    return i
        return cuda_index

    def _loop(self):
        while not self.stop_flag:
            try:
                if self.nv is not None:
                    self.sm.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                    self.mx.append(float(self.nv.nvmlDeviceGetMaxClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                    try:
                        m = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        m = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for name, bit in self.BITS.items():
                        if m & bit:
                            self.reasons.add(name)
                    time.sleep(0.0005)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
                r = [c.strip() for c in out.splitlines()[0].split(",")] if out else []
                if len(r) >= 9:
                    self.sm.append(float(r[1])); self.mx.append(float(r[2]))
                    for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                        if v.lower().startswith("active"):
                            self.reasons.add(name)
            except Exception:
                time.sleep(0.01)

    def start(self):
        self.th = threading.Thread(target=self._loop, daemon=True)
        self.th.start()

    def stop(self):
        self.stop_flag = True
        if self.th:
            self.th.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source}


# ---- CPU side: the reference's own renderer (oracle/_ref) or, if it could not be built, the C port -----------------
def cpu_scene():
    from oracle import bind as B
    path = mesh_path()
    if B.ref_available():
        return B, B.Scene(B.ref(), path, loader_threads=1 if MESH in ("Simple", "Cube") else min(8, host_threads()), rescale=RESCALE), "reference"
    if not B.port_available():
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], stdout=subprocess.DEVNULL)
    return B, B.Scene(B.port(), path, rescale=RESCALE), "port"


def cpu_frame_seconds(scene, kind, threads):
    """One full 4K frame on the host cores. reference: through start_render_from_camera / wait_… with `threads` pool
    threads (renderer.cpp:403-471); port: the oracle's row-parallel loop."""
    if kind == "reference":
        _, sec, _ = scene.render(W, H, spp=SPP, aa=AA, threads=threads)
        return sec
    t0 = time.perf_counter()
    if AA:
        scene.render_jitter(W, H, SPP, 0x853C49E6748FEA9B, threads=threads)
    else:
        scene.trace_primary(W, H, threads=threads)
    return time.perf_counter() - t0


def base_line(n_gpus, steps, warmup):
    return {"metric": METRIC, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": steps, "warmup": warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "name": CONFIG, "resolution": [W, H], "spp": SPP, "anti_aliasing": AA, "bounce_limit": 1, "tree": "SAH/300",
                       "mesh": "standin-v1 (Assets/Dragon.obj absent from the reference checkout)" if MESH == "standin" else "Assets/%s.obj" % MESH}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    _, scene, kind = cpu_scene()
    threads = host_threads()
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_frame_seconds(scene, kind, threads)
    secs = [cpu_frame_seconds(scene, kind, threads) for _ in range(args.steps)]
    total = sum(secs)
    val = W * H * SPP * len(secs) / total / 1e6
    line = base_line(args.gpus, args.steps, args.warmup)
    line.update({"impl": "reference", "value": val, "ms_per_step": 1e3 * total / len(secs),
                 "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind,
                                  "sample": "%d full frames (%d rays each) through the reference's tile thread pool" % (len(secs), W * H * SPP)},
                 "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    line["config"]["host_threads"] = threads
    emit(line)


# ---- our arm -----------------------------------------------------------------------------------------------------------
def ops_and_bytes_per_ray(c):
    """SURVEY §8(d) formula from the kernel's own work counters (one counted frame outside the timed region)."""
    rays = max(1, c["rays"])
    ops = (12 * c["box_tests"] + 14 * c["tri_tests"] + 10 * c["pass_det"] + 16 * c["pass_u"] + 6 * c["pass_v"]) / rays + 40
    byts = (32 * c["box_tests"] + 48 * c["tri_tests"]) / rays + 12
    return ops, byts


def load_traffic(key="dram_bytes_per_launch"):
    """dram bytes (or warp instructions) per launch of the render kernel from the committed ncu capture (profiles/), or None."""
    p = os.path.join(ROOT, "profiles", "render_kernel_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:
            return None
    return None


def pick_device(local_rank, world, visible):
    """GPU of a local rank. When the node shows at least twice as many GPUs as there are ranks, the ranks are spread evenly
    over them (cuda:LOCAL_RANK * (visible // world)): on the 8-GPU B200 boxes GPUs 0-3 and 4-7 reach host memory through
    two different paths (NVML reports one level for every pair, the copy rates do not: profiles/r01d_summary.md), and the
    end-to-end frame is bound by exactly that path. ATRAY_BENCH_DEVICES="0,1" (one entry per local rank) overrides."""
    env = os.environ.get("ATRAY_BENCH_DEVICES")
    if env:
        devs = [int(x) for x in env.split(",")]
        if len(devs) >= world and all(0 <= d < visible for d in devs):
            return devs[local_rank]
    if world > 1 and visible >= 2 * world:
        return local_rank * (visible // world)
    return local_rank


def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    visible_gpus = 1
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        visible_gpus = torch.cuda.device_count()
        local_rank = pick_device(local_rank, world, visible_gpus)
        torch.cuda.set_device(local_rank)
        dist_mod.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod
    if args.gpus != world:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 with torch.distributed.run)" % (args.gpus, world))

    import atray_b200 as A
    from atray_b200 import multi

    fp_mode = A.FP_FAST if args.fp == "fast" else A.FP_STRICT
    path = mesh_path() if rank == 0 else None
    if dist is not None:
        dist.barrier()
        path = mesh_path()
    t_setup = time.perf_counter()
    model = A.load_model_data(path, 1 if MESH in ("Simple", "Cube") else min(8, host_threads()))
    t_load = time.perf_counter()
    A.place_default(model, rescale=RESCALE)           # app.cpp:68-72 (no resize for the dragon; SURVEY 8d placement for the others)
    scene = A.Scene(model, division_method=A.DIV_SAH, max_no_faces_per_node=300)
    A.build_KD_tree(model, scene.division_method, scene.max_no_faces_per_node)   # (prep_scene builds only if the model has no tree)
    t_build = time.perf_counter()
    scene.ctx = A.Context(local_rank)                 # CUDA context creation: the bulk of the set-up on a fresh process
    t_ctx = time.perf_counter()
    A.prep_scene(scene, device=local_rank)            # flatten + upload once
    setup_s = time.perf_counter() - t_setup
    setup_parts = {"obj_load": t_load - t_setup, "place_and_tree_build": t_build - t_load, "cuda_context": t_ctx - t_build,
                   "flatten_and_upload": t_setup + setup_s - t_ctx}
    ctx = scene.ctx
    cam = A.set_camera(W, H, samples_per_pixel=SPP, anti_aliasing=AA)
    # GPU 0 gathers the BGRX bitmap (what the live view shows, north_star); t and primitive ids stay on the rank that
    # traced them: 4 instead of 12 bytes per pixel cross NVLink into GPU 0 (device-resident leg; the e2e leg writes the host
    # bitmap directly from every rank, see below)
    gather = multi.FrameGather(ctx, dist, rank, world, W, H, local=() if args.gather == "all" else ("t", "prim"))
    # N > 1: the per-frame view pass is shared — every rank computes 1/N of the slot ranges and stores them into all ranks'
    # tables over NVLink (atray_view_share_*, CUDA-IPC between the rank processes)
    share = multi.ViewShare(ctx, dist, rank, world, W, H) if (world > 1 and args.share and not args.no_share) else None
    opts = ctx.opts(fp_mode=fp_mode, rank=rank, world=world, band_rows=BAND_ROWS)
    my_rays = multi.rays_of_rank(W, H, SPP, BAND_ROWS, rank, world)
    flush_bytes = 256 << 20                            # > 126 MB L2
    d_flush = ctx.dev_alloc(flush_bytes)

    def flush_l2():
        # device-side write of a buffer larger than L2, on the context stream, outside the per-step event pair
        ctx.dev_memset(d_flush, 0, flush_bytes)

    def barrier():
        ctx.sync()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    def step_device():
        # every rank's frame starts when all ranks are ready (flag exchange queued on the streams, outside the step's event
        # pair): a step's duration is then the frame's, not the skew between the Python loops of the rank processes
        if share is not None:
            ctx.view_share_rendezvous()
        ctx.render_device(cam, opts, gather.planes["bgrx"], gather.planes["t"], gather.planes["prim"])

    # ---- HBM-resident leg: `value` -------------------------------------------------------------------------------------
    # clocks are sampled from the first warm-up frame to the last end-to-end frame (a timed leg lasts only milliseconds)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        flush_l2()
        step_device()
    barrier()
    wall0 = time.perf_counter()
    per_step_ms, render_ms, split_ms = [], [], []
    for _ in range(args.steps):
        flush_l2()
        step_device()
        per_step_ms.append(ctx.last_kernel_ms())       # waits for this step's stop event (device time on our stream)
        render_ms.append(ctx.last_render_kernel_ms())  # k_render_culled alone (event after the per-frame view pass)
        split_ms.append(ctx.last_view_split_ms())      # k_view_bounds | k_view_nodes
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - wall0)
    dev_ms = float(sum(per_step_ms))
    gather_check = None
    if dist is not None:
        import torch
        # a frame is complete when its slowest rank is: per step the maximum over the ranks, summed over the steps
        mine = torch.tensor(per_step_ms, dtype=torch.float64, device="cuda")
        allsteps = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allsteps, mine)
        steps_by_rank = torch.stack(allsteps).cpu().numpy()          # (world, K)
        dev_ms_max = float(steps_by_rank.max(axis=0).sum())
        per_rank_ms = [float(x) for x in steps_by_rank.mean(axis=1)]
        max_of_sums_ms = float(steps_by_rank.sum(axis=1).max())
        # what GPU 0 gathered (every rank's peer stores) is the frame one GPU renders alone
        if rank == 0:
            want = ("bgrx", "t", "prim") if args.gather == "all" else ("bgrx",)
            ref = ctx.render(cam, ctx.opts(fp_mode=fp_mode), want=want)
            gather_check = True
            for k in want:
                got = ctx.to_host(gather.planes[k], np.zeros((H, W), np.uint32))
                gather_check = gather_check and bool((got == np.ascontiguousarray(ref[k]).view(np.uint32)).all())
        barrier()
    else:
        dev_ms_max = dev_ms
        per_rank_ms = [dev_ms / args.steps]
        max_of_sums_ms = dev_ms
    total_rays = W * H * SPP * args.steps
    value = total_rays / (dev_ms_max * 1e-3) / 1e6

    # ---- end-to-end leg through the reference-facing call with the HOST bitmap ---------------------------------------------
    e2e = None
    e2e_check = None
    tex = A.PinnedArray((H, W), np.uint32) if rank == 0 else None
    if world == 1:
        info = A.RenderInfo(cam, scene, tex.array, opts=opts)
        for _ in range(2):
            A.start_render_from_camera(info)
            while A.wait_for_render_from_camera_to_finish(info, None, 33):
                pass
        ctx.sync()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            A.start_render_from_camera(info)            # kernel + D2H of the BGRX bitmap queued
            while A.wait_for_render_from_camera_to_finish(info, None, 33):
                pass
        e2e_s = time.perf_counter() - t0
    else:
        # every rank renders its bands on its own GPU and copies exactly those scanlines into ONE host bitmap in shared
        # pinned memory (atray_render_begin with a partition): N device→host copies in parallel, one per PCIe link
        name = "atray_bench_%s" % os.environ.get("MASTER_PORT", "0")
        shared = multi.SharedBitmap(name, (H, W), np.uint32, create=True) if rank == 0 else None
        shm_bar = multi.ShmBarrier(name + "_bar", rank, world, create=True) if rank == 0 else None
        barrier()
        if rank != 0:
            shared = multi.SharedBitmap(name, (H, W), np.uint32, create=False)
            shm_bar = multi.ShmBarrier(name + "_bar", rank, world, create=False)
        info = A.RenderInfo(cam, scene, shared.array, opts=opts)

        def e2e_step():
            A.start_render_from_camera(info)
            while A.wait_for_render_from_camera_to_finish(info, None, 33):
                pass
            shm_bar.wait()                              # the frame is complete when every rank's scanlines have landed
        for _ in range(2):
            e2e_step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        e2e_s = time.perf_counter() - t0
        if rank == 0:                                   # outside the timed region: the shared bitmap is the 1-GPU frame
            whole = A.RenderInfo(cam, scene, tex.array, opts=ctx.opts(fp_mode=fp_mode))
            A.start_render_from_camera(whole)
            while A.wait_for_render_from_camera_to_finish(whole, None, 33):
                pass
            e2e_check = bool((np.asarray(shared.array) == tex.array).all())
        barrier()
        shared.close()
        shm_bar.close()
    if dist is not None:
        import torch
        tt = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt.item())
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "warm-up + value leg + e2e leg"
    e2e = {"value": total_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 136,
           "d2h_bytes_per_step": W * H * 4, "ms_per_step": 1e3 * e2e_s / args.steps,
           "output": ("one host bitmap in shared pinned memory; every rank copies its own scanlines (parallel PCIe links); "
                      "equal to the 1-GPU frame: %s" % e2e_check) if world > 1 else "pinned host bitmap, slab-pipelined copy-out",
           "note": "host BGRX bitmap (Texture.bmb.buffer_memory, pinned) is the output; input per step is the camera/settings "
                   "struct passed as kernel parameters; the scene is uploaded once in prep_scene as the reference builds its tree once"}

    # ---- explanation: work counters, K0 peaks, roofline, CPU baseline (rank 0, outside the timed regions) ------------------
    line = None
    if rank == 0:
        one = ctx.opts(fp_mode=fp_mode, count_work=True, rank=0, world=1, band_rows=BAND_ROWS)
        ctx.render_device(cam, one, gather.planes["bgrx"], gather.planes["t"], gather.planes["prim"])
        ctx.sync()
        ctr = ctx.counters()
        ops_ray, bytes_ray = ops_and_bytes_per_ray(ctr)
        k0 = ctx.microbench()
        hbm_peak, peak_src = measured_peaks()
        # dominant kernel = k_render_culled: one launch per step per rank, timed alone with CUDA events on the library's stream
        # (after k_view_bounds + k_view_finish); per-launch algorithmic bytes = bytes/ray × rays of the launch
        launch_ms = float(sum(render_ms)) / args.steps
        launch_rays = my_rays
        achieved_gbs = bytes_ray * launch_rays / (launch_ms * 1e-3) / 1e9
        fp32_achieved = ops_ray * launch_rays / (launch_ms * 1e-3)
        bound_fp32 = k0["fp32_lane_ops_per_s"] / ops_ray / 1e6         # Mrays/s per GPU
        bound_hbm = hbm_peak * 1e9 / bytes_ray / 1e6
        per_gpu = value / world
        # The kernel is bound by instruction issue / dependent-operation latency, not by HBM (measured DRAM traffic is the output
        # planes, ~4 % of the HBM peak): `issue` is the fraction of the SMs' issue slots the kernel's warp instructions fill
        # (warp instructions per launch from the committed ncu capture of the same workload, profiles/render_kernel_traffic.json).
        warp_inst = load_traffic("warp_instructions_per_launch")
        sm_hz = 1e6 * (clocks["sm_mhz"] if clocks and clocks.get("sm_mhz") else 1965.0)
        issue = None
        if warp_inst and world == 1:
            issue = {"warp_instructions_per_launch": warp_inst, "issue_slots_per_s": 148 * 4 * sm_hz,
                     "frac": warp_inst / (148 * 4 * sm_hz * launch_ms * 1e-3), "source": "profiles/render_kernel_traffic.json (ncu --set full)"}
        roofline = {"bound": "hbm", "bound_note": "nominal: the contract's HBM line; the kernel is issue/latency-bound, see `issue` and fp32",
                    "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": achieved_gbs / hbm_peak,
                    "traffic": load_traffic(), "peak_source": peak_src, "kernel": "k_render_culled<STRICT>" if args.fp == "strict" else "k_render_culled<FAST>",
                    "kernel_ms_per_launch": launch_ms, "kernel_share_of_step": launch_ms / (dev_ms / args.steps),
                    "algorithmic_bytes_per_ray": bytes_ray, "algorithmic_fp32_ops_per_ray": ops_ray,
                    "issue": issue,
                    "fp32": {"achieved_lane_ops_per_s": fp32_achieved, "peak_lane_ops_per_s": k0["fp32_lane_ops_per_s"],
                             "frac": fp32_achieved / k0["fp32_lane_ops_per_s"], "peak_source": "K0 FFMA microbenchmark, this run"},
                    "l2_read_gbs_measured": k0["l2_read_gbs"],
                    "north_star_roofline_mrays_per_gpu": {"fp32_issue_bound": bound_fp32, "hbm_bytes_bound": bound_hbm,
                                                          "slower": min(bound_fp32, bound_hbm),
                                                          "achieved_frac_of_slower": per_gpu / min(bound_fp32, bound_hbm)},
                    "note": "achieved = algorithmic bytes per ray (32 B per box test + 48 B per triangle test actually made, "
                            "+ 12 B output; counters of a separate counted frame) x rays / render-kernel time. The 10 MB scene is "
                            "L2/L1-resident, so DRAM traffic is the output planes; the kernel is bound by dependent-operation "
                            "latency per warp, not by HBM (see issue.frac, fp32.frac, l2_read_gbs_measured and profiles/)"}
        cpu_baseline = None
        if world == 1 and not args.no_cpu_baseline:
            _, cscene, kind = cpu_scene()
            threads = host_threads()
            cpu_frame_seconds(cscene, kind, threads)
            secs = [cpu_frame_seconds(cscene, kind, threads) for _ in range(3)]
            cpu_baseline = {"value": W * H * SPP / min(secs) / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                            "sample": "3 full frames (%d primary rays each) after 1 warm-up, best of 3, "
                                      "reference tile thread pool with %d threads" % (W * H * SPP, threads),
                            "seconds": secs}
        line = base_line(args.gpus, args.steps, args.warmup)
        line.update({"impl": "ours", "value": value, "ms_per_step": dev_ms_max / args.steps, "e2e": e2e, "roofline": roofline,
                     "cpu_baseline": cpu_baseline, "clocks": clocks,
                     # timed `value` region: k_view_bounds + k_view_finish + k_render_culled per step and rank
                     "gpu_launches": (3 if share is None else 6) * args.steps * world,   # k_view_bounds, k_view_nodes, k_render_culled (+ k_view_stamp and the two flag kernels of the rendezvous)
                     "wall_ms_bracket": wall_ms, "per_step_ms_rank0": per_step_ms, "ms_per_step_by_rank": per_rank_ms,
                     "ms_per_step_max_of_rank_sums": max_of_sums_ms / args.steps, "gathered_frame_equals_one_gpu_frame": gather_check,
                     "kernel_ms_rank0": {"k_view_bounds": float(np.mean([a for a, _ in split_ms])), "k_view_nodes": float(np.mean([b for _, b in split_ms])),
                                         "k_render_culled": float(np.mean(render_ms))},
                     "counters": ctr, "setup_s": setup_s, "setup_parts_s": setup_parts})
        line["config"].update({"fp_mode": args.fp, "band_rows": BAND_ROWS, "parallelism": "bands%d" % world,
                               "l2": "flushed between timed steps (256 MiB device memset outside the per-step event pair)",
                               "timing": "CUDA events on the library's stream around each frame's launch; per step the MAX over the ranks, summed over the K steps"
                                         + ("; every step starts after an on-stream flag rendezvous of the ranks" if share is not None else ""),
                               "view_pass": ("shared: each rank computes 1/%d of the slot ranges and peer-stores them into all ranks' tables (NVLink, CUDA IPC)" % world) if share is not None else "computed by every rank",
                               "devices": [pick_device(r, world, visible_gpus) for r in range(world)] if world > 1 else [local_rank],
                               "gather": ("none (1 GPU)" if world == 1 else
                                          "all three planes (12 B/px, SURVEY 8e): direct peer stores into GPU 0's frame through CUDA IPC" if args.gather == "all" else
                                          "BGRX bitmap: direct peer stores into GPU 0's frame through CUDA IPC; t/prim planes stay on the tracing rank")})
    barrier()
    if tex is not None:
        tex.free()
    ctx.dev_free(d_flush)
    if share is not None:
        ctx.view_share_check()
        share.close()
    gather.close()
    if dist is not None:
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


# ---- C5: 64 incoherent rays per pixel index (BASELINE.json configs[4]) ------------------------------------------------------
C5_SEED = 0x853C49E6748FEA9B


def c5_cpu_sample(n_px, threads):
    """The reference's own traversal + triangle test (oracle/_ref: get_ray_kd_tree_intersection through get_intersection_data)
    on the C5 rays of `n_px` pixel indices from the middle of the frame — rays written out by the oracle's restatement of the
    stream, which is not timed. Returns (Mrays/s, kind, rays, seconds)."""
    from oracle import bind as B
    path = mesh_path()
    if not B.port_available():
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], stdout=subprocess.DEVNULL)
    port = B.Scene(B.port(), path, rescale=RESCALE)
    p0 = W * H // 2
    _, _, _, rays = port.trace_c5(p0, p0 + n_px, C5_SPP, C5_SEED, threads=threads, want_rays=True)
    if B.ref_available():
        sc, kind = B.Scene(B.ref(), path, loader_threads=min(8, threads), rescale=RESCALE), "reference"
    else:
        sc, kind = port, "port"
    sc.trace_rays(rays[:4096], threads=threads)
    t0 = time.perf_counter()
    sc.trace_rays(rays, threads=threads)
    sec = time.perf_counter() - t0
    return rays.shape[0] / sec / 1e6, kind, int(rays.shape[0]), sec


def run_reference_c5(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = host_threads()
    vals = [c5_cpu_sample(16384, threads) for _ in range(max(1, min(args.steps, 3)))]
    val = float(np.mean([v[0] for v in vals]))
    line = base_line(args.gpus, args.steps, args.warmup)
    line.update({"impl": "reference", "value": val, "ms_per_step": 1e3 * W * H * C5_SPP / (val * 1e6),
                 "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": vals[0][1],
                                  "sample": "%d rays (the 64 rays of 16384 pixel indices from the middle of the frame) per step through the "
                                            "reference's traversal + triangle test, %d threads; ms_per_step = the full 531 M-ray step at that rate"
                                            % (vals[0][2], threads)},
                 "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})
    line["config"]["host_threads"] = threads
    emit(line)


def run_ours_c5(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local_rank)
        dist_mod.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod
    if args.gpus != world:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 with torch.distributed.run)" % (args.gpus, world))
    import atray_b200 as A
    from atray_b200 import multi
    fp_mode = A.FP_FAST if args.fp == "fast" else A.FP_STRICT
    path = mesh_path() if rank == 0 else None
    if dist is not None:
        dist.barrier()
        path = mesh_path()
    t_setup = time.perf_counter()
    model = A.load_model_data(path, min(8, host_threads()))
    A.place_default(model, rescale=RESCALE)
    scene = A.Scene(model, division_method=A.DIV_SAH, max_no_faces_per_node=300)
    A.prep_scene(scene, device=local_rank)
    setup_s = time.perf_counter() - t_setup
    ctx = scene.ctx
    n_px = W * H
    # the pixel-colour plane lives on GPU 0; every rank stores the colours of its chunks into it (peer stores over NVLink)
    gather = multi.FrameGather(ctx, dist, rank, world, W, H, planes=("bgrx",))
    d_bgrx = gather.planes["bgrx"]
    my_px = sum(min(C5_CHUNK, n_px - k * C5_CHUNK) for k in range(rank, (n_px + C5_CHUNK - 1) // C5_CHUNK, world))

    def barrier():
        ctx.sync()
        if dist is not None:
            import torch
            torch.cuda.synchronize()
            dist.barrier()

    def step():
        ctx.trace_incoherent_device(0, n_px, C5_SPP, d_bgrx, 0, 0, seed=C5_SEED, rank=rank, world=world, chunk_px=C5_CHUNK, fp_mode=fp_mode)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    per_step_ms = []
    d_flush = ctx.dev_alloc(256 << 20)
    for _ in range(args.steps):
        ctx.dev_memset(d_flush, 0, 256 << 20)     # L2 flushed between the timed steps (outside the step's event pair)
        step()
        per_step_ms.append(ctx.last_kernel_ms())
    ctx.dev_free(d_flush)
    barrier()
    if dist is not None:
        import torch
        mine = torch.tensor(per_step_ms, dtype=torch.float64, device="cuda")
        allsteps = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allsteps, mine)
        by_rank = torch.stack(allsteps).cpu().numpy()
    else:
        by_rank = np.array([per_step_ms])
    dev_ms_max = float(by_rank.max(axis=0).sum())
    total_rays = n_px * C5_SPP * args.steps
    value = total_rays / (dev_ms_max * 1e-3) / 1e6
    # end to end: the same step through the host-facing call — every rank traces its chunks, then the colour plane (on GPU 0)
    # is copied to a pinned host bitmap
    tex = A.PinnedArray((H, W), np.uint32) if rank == 0 else None
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
        barrier()
        if rank == 0:
            ctx.to_host(d_bgrx, tex.array)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    line = None
    if rank == 0:
        # parity inside the bench: the colours of a sub-range against the oracle (outside the timed regions)
        check = None
        try:
            from oracle import bind as B
            o = B.Scene(B.port(), path, rescale=RESCALE)
            _, _, ref_px = o.trace_c5(n_px // 2, n_px // 2 + 2048, C5_SPP, C5_SEED, threads=host_threads())
            check = bool((tex.array.reshape(-1)[n_px // 2:n_px // 2 + 2048] == ref_px).all())
        except Exception as e:   # the checker is optional here (tests/ pins the path)
            check = "not checked: %s" % e
        sub = 65536
        d_tmp = ctx.dev_alloc(sub * 4)
        ctx.trace_incoherent_device(n_px // 2, n_px // 2 + sub, C5_SPP, d_tmp, 0, 0, seed=C5_SEED, fp_mode=fp_mode, count_work=True)
        ctx.sync()
        ctr = ctx.counters()
        ctx.dev_free(d_tmp)
        ops_ray, bytes_ray = ops_and_bytes_per_ray(ctr)
        k0 = ctx.microbench()
        hbm_peak, peak_src = measured_peaks()
        launch_ms = float(np.mean(per_step_ms))
        achieved_gbs = bytes_ray * my_px * C5_SPP / (launch_ms * 1e-3) / 1e9
        fp32_achieved = ops_ray * my_px * C5_SPP / (launch_ms * 1e-3)
        roofline = {"bound": "hbm", "bound_note": "nominal: the scene (12 MB) is L2-resident; the kernel is bound by divergent instruction issue",
                    "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": achieved_gbs / hbm_peak, "traffic": None,
                    "peak_source": peak_src, "kernel": "k_c5<STRICT>" if args.fp == "strict" else "k_c5<FAST>", "kernel_ms_per_launch": launch_ms,
                    "kernel_share_of_step": 1.0, "algorithmic_bytes_per_ray": bytes_ray, "algorithmic_fp32_ops_per_ray": ops_ray,
                    "fp32": {"achieved_lane_ops_per_s": fp32_achieved, "peak_lane_ops_per_s": k0["fp32_lane_ops_per_s"],
                             "frac": fp32_achieved / k0["fp32_lane_ops_per_s"], "peak_source": "K0 FFMA microbenchmark, this run"},
                    "l2_read_gbs_measured": k0["l2_read_gbs"],
                    "note": "counters of a 65536-pixel sub-range (4.2 M rays): 32 B per box test + 48 B per triangle test + 4 B/64 output per ray"}
        cpu_baseline = None
        if world == 1 and not args.no_cpu_baseline:
            v, kind, nr, sec = c5_cpu_sample(16384, host_threads())
            cpu_baseline = {"value": v, "unit": "Mrays/s", "cores": host_threads(), "kind": kind,
                            "sample": "%d rays (64 per pixel index, 16384 pixel indices) in %.1f s" % (nr, sec)}
        line = base_line(args.gpus, args.steps, args.warmup)
        line.update({"impl": "ours", "value": value, "ms_per_step": dev_ms_max / args.steps, "scaling": "strong",
                     "e2e": {"value": total_rays / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": n_px * 4,
                             "ms_per_step": 1e3 * e2e_s / args.steps,
                             "note": "rays are generated on the device from the per-ray stream (input = the job description); output = the "
                                     "pixel-colour plane copied from GPU 0 to a pinned host bitmap every step"},
                     "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": clocks, "gpu_launches": args.steps * world,
                     "ms_per_step_by_rank": [float(x) for x in by_rank.mean(axis=1)], "colours_equal_oracle_on_2048_pixels": check,
                     "counters_sub_range": ctr, "setup_s": setup_s})
        line["config"].update({"fp_mode": args.fp, "parallelism": "chunks%d" % world, "chunk_px": C5_CHUNK, "rays_per_step": n_px * C5_SPP,
                               "timing": "CUDA events on the library's stream around each step's launch; per step the MAX over the ranks",
                               "l2": "flushed between timed steps (256 MiB device memset outside the per-step event pair)",
                               "gather": "pixel colours: peer stores into GPU 0's plane (CUDA IPC)" if world > 1 else "none (1 GPU)"})
    barrier()
    if tex is not None:
        tex.free()
    gather.close()
    if dist is not None:
        dist.destroy_process_group()
    if rank == 0:
        emit(line)


_RESULT_FD = None


def emit(line):
    """The result line goes to the process's ORIGINAL stdout (see main)."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    # The contract is ONE JSON line on stdout, but native libraries write there too (NCCL prints its version banner with
    # printf when NCCL_DEBUG is set in the environment): keep the original stdout for the result line and point fd 1 at
    # stderr for everything else.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--fp", choices=["strict", "fast"], default="strict")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config", choices=sorted(CONFIGS), default="c4", help="BASELINE.json configs[0..4]; c4 is the headline")
    ap.add_argument("--gather", choices=("bgrx", "all"), default="bgrx",
                    help="N > 1, device-resident leg: what GPU 0 gathers over NVLink — the bitmap (4 B/px, what the live view shows) or "
                         "bitmap + hit distance + primitive id (12 B/px)")
    ap.add_argument("--share", action="store_true",
                    help="N > 1: share the per-frame view pass between the ranks over NVLink (atray_view_share_*); measured slower than "
                         "every rank computing the pass itself at N = 2, 4 and 8 (profiles/r02i, r02j), so it is opt-in")
    ap.add_argument("--no-share", action="store_true", help="(the default since r02i) every rank computes the whole per-frame view pass itself")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    apply_config(args.config)
    if args.config == "c5":
        if args.steps == 100:
            args.steps = 3   # a step is 531 M rays (seconds on one GPU): the default run stays within minutes
        if args.impl == "reference":
            run_reference_c5(args)
        else:
            run_ours_c5(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
