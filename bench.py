#!/usr/bin/env python
"""bench.py — fwd+bwd Mpix/s of the rasterizer hot path (BASELINE.json metric) on synthetic scenes.

    python bench.py --gpus N --steps K --warmup W            # B200 arm (this repo's CUDA path)
    python bench.py --impl reference --gpus N --steps K ...   # CPU arm: oracle/_cpu (C/OpenMP port) on host cores

A "step" = every rank renders `views_per_rank` 1920x1080 views of the replicated 1 M-Gaussian SH-3 scene
forward + backward (gradients w.r.t. all five activated tensors accumulate), and, for N>1, the sum-all-reduce of
the gradient (five NCCL collectives below 4 ranks; from 4 ranks on this library's own multimem kernel over a
symmetric buffer the .grad tensors live in — GradientArena).  value = all ranks' pixels / max-over-ranks device time.
Prints ONE JSON line on rank 0 (keys: see DESIGN.md "Measurement"): value / value_one_stream / e2e / roofline /
cpu_baseline / reference_trainer (the reference's unmodified _training_step on this backend) / f1_store_path /
train_step / multi_gpu diagnostics / oracle_pin.
"""
from __future__ import annotations

import argparse
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (N gaussians, W, H, sh_degree)
    "A": (1_000_000, 1920, 1080, 3),   # BASELINE.json configs[1] — the configuration the metric is quoted on
    "C": (6_000_000, 3840, 2160, 3),   # configs[3]
    "B": (3_000_000, 1920, 1080, 3),   # configs[2] (per-rank view shard of the 64-camera set)
    "tiny": (20_000, 320, 192, 3),
}
METRIC = "fwd+bwd Mpix/s (1M gaussians SH3, 1920x1080)"  # BASELINE.json's metric, quoted on config A


def metric_name(config: str) -> str:
    """BASELINE.json's metric string for config A; the same quantity labelled with the actual sizes otherwise."""
    if config == "A":
        return METRIC
    N, W, H, deg = CONFIGS[config]
    return f"fwd+bwd Mpix/s ({N / 1e6:g}M gaussians SH{deg}, {W}x{H})"
LAST_PAIRS = None  # (pixel x Gaussian pairs visited, composited) by the CPU port's last forward blend


def config_dict(name: str, V: int, n_vis, M, world: int = 1, total_views: int = 0) -> dict:
    """The `config` object — identical on the B200 arm and the `--impl reference` arm for the same flags."""
    N, W, H, deg = CONFIGS[name]
    return {"workload": f"{name}: {N} gaussians SH{deg}, {W}x{H}, {V} view(s)/rank/step fwd+bwd, cameras = circle_cameras(64)",
            "N": N, "N_visible": n_vis, "M_isects": M, "views_per_rank": V,
            "l2": "inputs larger than L2 (params 236 B/Gaussian + keys/records per view > 126 MB); no explicit flush"}


def view_index(step: int, v: int, V: int, rank: int = 0, world: int = 1) -> int:
    """Camera of view `v` of step `step` on `rank` (same sequence on both arms)."""
    return (step * world * V + rank * V + v) % 64


def gsplat_probe() -> str:
    """is the reference's real arithmetic provider (gsplat==1.5.3, requirements.txt:9) importable here?"""
    import importlib.util
    try:
        if importlib.util.find_spec("gsplat") is None:
            return "gsplat unavailable on this box"
        import gsplat
        if getattr(gsplat, "__version__", "") == "b200-shim":
            return "gsplat unavailable on this box"
        return f"gsplat {gsplat.__version__} importable (pin: tests/test_gpu_gsplat_pin.py)"
    except Exception as e:
        return f"gsplat unavailable on this box (import failed: {type(e).__name__})"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.samples = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[1]))
                mx.append(float(s[2]))
                for n, v in zip(names, s[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class StageTimer:
    """CUDA-event recorder for the per-stage hooks in the package (events on the launching stream)."""

    def __init__(self, torch):
        self.torch = torch
        self.open = {}
        self.pairs = {}
        self.enabled = False

    def begin(self, name):
        if not self.enabled:
            return
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.open[name] = e

    def end(self, name):
        if not self.enabled:
            return
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        self.pairs.setdefault(name, []).append((self.open.pop(name), e))

    def totals_ms(self):
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self.pairs.items()}

    def counts(self):
        return {k: len(v) for k, v in self.pairs.items()}


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle's C/OpenMP port (oracle/cpu_splat.c -> oracle/_cpu/libcpusplat.so)
# ---------------------------------------------------------------------------------------------------
def cpu_port_time(N, W, H, deg, steps, warmup, seed=0, views_per_step=1):
    """Times fwd+bwd of `views_per_step` views per step on the host cores with the C port of the oracle (same scene,
    same camera sequence as the B200 arm).  Returns (seconds_per_step, cores, M of camera 0, N_visible of camera 0)."""
    from oracle import cpu_port
    import torch
    from oracle import torch_oracle as orc

    cpu_port.set_threads(os.cpu_count() or 1)  # all host cores, whatever OMP_NUM_THREADS the launcher exported
    sc = orc.synthetic_scene(N, W, H, seed=seed)
    act = orc.activate(sc)
    cams = orc.circle_cameras(64)
    g = torch.Generator().manual_seed(1)
    v_render = torch.randn(H, W, 3, generator=g)
    fw0 = cpu_port.forward(act, cams[0], sc["K"], W, H, deg, bg=torch.zeros(3))
    M0, nvis0 = fw0["M"], int((fw0["radii"] > 0).all(-1).sum())
    global LAST_PAIRS
    LAST_PAIRS = cpu_port.pairs_evaluated()
    del fw0
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        for v in range(views_per_step):
            cpu_port.forward_backward(act, cams[view_index(i, v, views_per_step)], sc["K"], W, H, deg, v_render)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return sum(times) / len(times), cpu_port.num_threads(), M0, nvis0


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    N, W, H, deg = CONFIGS[args.config]
    V = args.views_per_rank if not args.total_views else max(1, args.total_views // max(1, args.gpus))
    # the SAME workload as the B200 arm — full N, full resolution, the same cameras, V views per step, exactly K timed steps
    # after W warm-up steps — unless that would take more than a few minutes on the host cores (config C / B at size):
    # then the step count is bounded (one view of config A is ~0.3-0.6 s on 16-32 threads)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    budget_views = 160 if args.config in ("A", "tiny") else 24
    if (steps + warm) * V > budget_views:
        steps = max(1, budget_views // V - min(warm, 1))
        warm = min(warm, 1)
    try:
        sec, cores, M, nvis = cpu_port_time(N, W, H, deg, steps, warm, views_per_step=V)
    except Exception as e:  # the port is part of the repo; failing to run it is an error worth seeing
        print(json.dumps({"impl": "reference", "unavailable": f"oracle C port failed: {e!r}"[:200]}))
        return 0
    mpix = V * W * H / sec / 1e6
    line = {
        "impl": "reference", "metric": metric_name(args.config), "value": mpix, "unit": "Mpix/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong" if args.total_views else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.config, V, nvis, M, args.gpus, args.total_views),
        "cpu_baseline": {"value": mpix, "unit": "Mpix/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} steps x {V} full views (N={N}, {W}x{H}, M={M} at camera 0) on {cores} host threads, "
                                   "oracle/cpu_splat.c (the reference has no CPU implementation of this path: "
                                   "src/backends/pytorch_backend.py:1-5 raises)"},
        "e2e": {"value": mpix, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "oracle_pin": gsplat_probe(),
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    orc = importlib.import_module("3d-gaussian-splatting_b200.synthetic")  # workload generator; oracle/ is not imported on this arm

    N, W, H, deg = CONFIGS[args.config]
    V = args.views_per_rank
    if args.total_views:  # strong scaling (config B: the 64-camera set split over the ranks)
        V = max(1, args.total_views // world)
    sc = orc.synthetic_scene(N, W, H, seed=0)
    act = orc.activate(sc)
    cams = orc.circle_cameras(64)
    leaves = {k: act[k].to(dev).clone().requires_grad_(True) for k in ("means3D", "scales", "rotations", "opacities", "shs")}
    K = sc["K"].to(dev)
    backend = pkg.B200Backend(K)
    bg = torch.zeros(3, device=dev)
    gen = torch.Generator().manual_seed(1 + rank)
    v_render = torch.randn(H, W, 3, generator=gen).to(dev)
    gt_host = torch.rand(H, W, 3, generator=gen).pin_memory()
    pose_host = [cams[i].clone().pin_memory() for i in range(64)]
    poses_dev = [cams[i].to(dev) for i in range(64)]
    params_list = [leaves[k] for k in ("means3D", "scales", "rotations", "opacities", "shs")]
    flat = None

    # views of a step alternate over `--view-streams` CUDA streams (streams.py): the short latency-bound kernels and the
    # drain phase of one view run under the long blend kernels of the other
    vstreams = pkg.ViewStreams(dev, args.view_streams, params_list)

    # Which camera this rank renders as view v of a step.  One GPU: the plain sequence.  N GPUs: the SAME global batch of
    # N*V cameras per step as the plain sequence, re-dealt to the ranks by `balance_views` on each camera's intersection
    # count M (measured once below): a step ends at the slowest rank, and cameras differ by ~10 % in M.
    assignment = {}
    cam_cost = None

    def my_view(step, v):
        if world == 1 or cam_cost is None or args.no_balance:
            return view_index(step, v, V, rank, world)
        period = 64 // math.gcd(64, world * V)
        key = step % period
        if key not in assignment:
            batch = [(key * world * V + j) % 64 for j in range(world * V)]
            parts = pkg.balance_views([cam_cost[c] for c in batch], world, V)
            assignment[key] = [[batch[i] for i in part] for part in parts]
        return assignment[key][rank][v]

    if world > 1 and not args.no_balance:
        with torch.no_grad():  # per-camera M, identical on every rank (same scene, same cameras)
            cam_cost = []
            for ci in range(64):
                camb = pkg.make_cam(poses_dev[ci], K, W, H)
                cam_cost.append(pkg.project_and_bin(leaves["means3D"], leaves["rotations"], leaves["scales"],
                                                    leaves["opacities"].squeeze(-1).contiguous(), leaves["shs"], deg, camb, W, H)["M"])
    probe = {"on": False, "render": [], "reduce": []}  # CUDA events of the multi-GPU diagnostic pass (not the headline)

    # N > 1: the five gradient tensors are views of ONE flat buffer (distributed.FlatGradients), reduced by one collective
    # (opt-in: measured at N=2 it costs more — the memset and the first view's in-place add, 0.2 ms — than the one collective
    # saves, 0.12 ms: 5.42 -> 5.56 ms per step)
    # --flat-grads sym: the flat buffer in symmetric memory, reduced by the library's multimem kernel (0.55 ms against NCCL's
    # 0.64 ms for one / 0.93 ms for five collectives at 8 ranks; slower than NCCL at 2 ranks) — the default from 8 ranks on
    # Measured (profiles/r02/bench_A_n*_{arena,off}.json): 'sym' / 'nccl' pay a memset and an extra in-place add of the first
    # view per step (FlatGradients), which costs the device leg more than the single collective saves; 'arena' has neither and
    # wins from 4 ranks on — value leg 5.63 -> 5.57 ms at 4 ranks, 5.73 -> 5.47 ms at 8; e2e leg 6.40 -> 6.25 and 7.18 -> 6.79 ms
    # — and loses at 2 (5.40 -> 5.55 ms: with two members the switch round trip of multimem buys nothing).
    # 'arena' (distributed.GradientArena) gets the same single own collective WITHOUT that bookkeeping: the first backward of a
    # step writes its gradients straight into slices of the symmetric buffer and autograd adopts them as .grad.
    default_mode = "arena" if world >= 4 else "off"
    fg_mode = args.flat_grads or default_mode
    fg_mode_e2e = args.flat_grads or default_mode
    flat_by_mode = {}
    arena_params = [leaves["means3D"], leaves["rotations"], leaves["scales"], leaves["opacities"], leaves["shs"]]

    class _Arena:  # FlatGradients' two-call interface over GradientArena
        def __init__(self):
            self.arena = pkg.GradientArena([leaves["means3D"], leaves["rotations"], leaves["scales"],
                                            leaves["opacities"].squeeze(-1), leaves["shs"]])
            self.hits = self.steps = 0

        def begin_step(self):
            for p in params_list:
                p.grad = None
            pkg.set_gradient_arena(self.arena)
            self.arena.begin_step()

        def allreduce(self, n_views):
            self.hits += int(self.arena.allreduce(arena_params, n_views))
            self.steps += 1
            pkg.set_gradient_arena(None)

    def make_flat(mode):
        if world == 1 or mode == "off":
            return None, mode
        if mode not in flat_by_mode:
            try:
                flat_by_mode[mode] = _Arena() if mode == "arena" else pkg.FlatGradients(params_list, symmetric=(mode == "sym"))
            except Exception as e:  # no multicast on this fabric: stay on NCCL's five collectives, and say so
                return None, f"off (symmetric memory unavailable: {type(e).__name__})"
        return flat_by_mode[mode], mode

    flat_grads, fg_mode = make_flat(fg_mode)
    flat_grads_e2e, fg_mode_e2e = make_flat(fg_mode_e2e)

    def step_device(step):
        """inputs resident in HBM: pose + cotangent on device."""
        nonlocal flat
        if flat_grads is not None:
            flat_grads.begin_step()
        else:
            for p in params_list:
                p.grad = None
        if probe["on"]:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            ev[0].record()
        vstreams.begin()
        if args.batched_backward:  # the reference's order (trainer.py:458-506): every view rendered, then the backward passes
            outs = []
            for v in range(V):
                with vstreams.view(v):
                    vp = {"world_view_transform": poses_dev[my_view(step, v)], "image_width": W, "image_height": H}
                    outs.append(backend.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda",
                                                          sh_degree=deg)["render"])
            for v in reversed(range(V)):
                with vstreams.view(v):
                    outs[v].backward(v_render)
            del outs
        else:
            for v in range(V):
                with vstreams.view(v):
                    vp = {"world_view_transform": poses_dev[my_view(step, v)], "image_width": W, "image_height": H}
                    res = backend.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
                    res["render"].backward(v_render)
        vstreams.join()
        if probe["on"]:
            ev[1].record()
        if flat_grads is not None:
            flat_grads.allreduce(world * V)
        elif world > 1:
            flat = pkg.allreduce_gradients(params_list, world * V, flat=flat)
        if probe["on"]:
            ev[2].record()
            probe["render"].append((ev[0], ev[1]))
            probe["reduce"].append((ev[1], ev[2]))

    # e2e: every view's pose and ground-truth frame come from pinned HOST memory.  The upload of view k+1 runs
    # on a copy stream while view k renders (double-buffered device slots, events both ways), as a training loop
    # with a prefetching loader does; each view's copy is issued inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    gt_slot = [torch.empty(H, W, 3, device=dev) for _ in range(2)]
    pose_slot = [torch.empty(4, 4, device=dev) for _ in range(2)]
    ready_ev = [torch.cuda.Event() for _ in range(2)]
    free_ev = [torch.cuda.Event() for _ in range(2)]
    e2e_state = {"k": 0, "primed": False}

    def upload(slot, view_idx):
        copy_stream.wait_event(free_ev[slot])  # the last reader of this slot (view k-2's loss/backward) is done
        with torch.cuda.stream(copy_stream):
            gt_slot[slot].copy_(gt_host, non_blocking=True)
            pose_slot[slot].copy_(pose_host[view_idx], non_blocking=True)
            ready_ev[slot].record(copy_stream)

    loss_pinned = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    e2e_mode = {"async": False, "last": None}

    def step_e2e(step):
        """through the public backend call with HOST buffers: pose + ground-truth frame uploaded from pinned
        memory every view, L1 + D-SSIM loss, backward, loss read back (what trainer.py:270-281,485-497 does per view).
        e2e_mode["async"]: the step's loss is still copied to the host every step, but into pinned memory without blocking;
        the host consumes it one step later (how a loop that only LOGS the loss would read it) — reported separately."""
        nonlocal flat
        if flat_grads_e2e is not None:
            flat_grads_e2e.begin_step()
        else:
            for p in params_list:
                p.grad = None
        losses = []
        vstreams.begin()
        for v in range(V):
            with vstreams.view(v):
                cur = torch.cuda.current_stream()
                k = e2e_state["k"]
                slot = k & 1
                if not e2e_state["primed"]:
                    for sl in range(2):
                        free_ev[sl].record(cur)
                    upload(slot, my_view(step, v))
                    e2e_state["primed"] = True
                nv, ns = (v + 1, step) if v + 1 < V else (0, step + 1)
                upload(slot ^ 1, my_view(ns, nv))  # prefetch the next view while this one renders
                cur.wait_event(ready_ev[slot])
                vp = {"world_view_transform": pose_slot[slot], "image_width": W, "image_height": H}
                res = backend.render_with_depth(leaves, vp, bg_color=bg, render_mode="RGB", device="cuda", sh_degree=deg)
                # the reference's per-view loss (trainer.py:485-492: clamp, L1 + 0.2 D-SSIM), value and gradient in one kernel
                loss = pkg.l1_ssim_loss(res["render"], gt_slot[slot], 0.2)["loss"]
                loss.backward()
                free_ev[slot].record(cur)
                losses.append(loss.detach())
                e2e_state["k"] = k + 1
        vstreams.join()
        if flat_grads_e2e is not None:
            flat_grads_e2e.allreduce(world * V)
        elif world > 1:
            flat = pkg.allreduce_gradients(params_list, world * V, flat=flat)
        total = torch.stack(losses).sum()
        if not e2e_mode["async"]:
            return float(total.item())  # D2H read of the step's result, blocking (the headline e2e)
        slot = step & 1
        loss_pinned[slot].copy_(total.reshape(1), non_blocking=True)
        loss_ev[slot].record()
        prev, e2e_mode["last"] = e2e_mode["last"], slot
        if prev is not None:
            loss_ev[prev].synchronize()
            return float(loss_pinned[prev].item())  # the PREVIOUS step's loss: already on the host
        return 0.0

    # f1: the same step from the RAW parameter store — activations fused into the kernels, gradients
    # accumulated in-kernel into the flat buffer that the all-reduce then takes as is
    store = pkg.GaussianStore.from_tensors(sc["xyz"].to(dev), sc["features_dc"].to(dev), sc["features_rest"].to(dev),
                                           sc["scaling"].to(dev), sc["rotation"].to(dev), sc["opacity"].to(dev))

    def step_store(step):
        store.zero_grad(lazy=True)  # the first backward writes the flat gradient buffer (no memset, no read-modify-write)
        vstreams.begin()
        for v in range(V):
            with vstreams.view(v):
                vp = {"world_view_transform": poses_dev[my_view(step, v)], "image_width": W, "image_height": H}
                res = store.render(K, vp, bg_color=bg, render_mode="RGB", sh_degree=deg)
                res["render"].backward(v_render)
        vstreams.join()
        if world > 1:
            pkg.allreduce_flat(store.flat_grad, 1.0 / (world * V))

    # f1+f3+f4: one optimiser iteration as trainer.py:447-519 runs it, on the widened path — V views from pinned host
    # memory (prefetched), GaussianStore.render, fused L1+SSIM loss (mean over views), backward with in-kernel gradient
    # accumulation, [all-reduce], one-kernel Adam, densification statistic; the loss is read back once per step
    ref_lrs = {"xyz": 1.6e-4, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
    # trainer.py:854-859 caps the log-scales every iteration at log(scene_extent * scale_clamp_ratio): folded into Adam
    adam = pkg.StoreAdam(store, ref_lrs, scale_ceiling=math.log(10.0 * 0.2 + 1e-9))
    train_state = {"k": 0, "primed": False}
    # N > 1: the optimiser of the train leg is the fused multimem kernel (gradient reduce-scatter + Adam on the rank's slice +
    # parameter all-gather, sharded moments) on a store in symmetric memory — when the fabric has multicast; else NCCL + Adam
    train_store, train_adam, train_mode = store, adam, ("single GPU" if world == 1 else "NCCL all-reduce + Adam on every rank")
    if world > 1 and not args.no_fused_optimizer:
        try:
            train_store = pkg.GaussianStore.from_tensors(sc["xyz"].to(dev), sc["features_dc"].to(dev), sc["features_rest"].to(dev),
                                                         sc["scaling"].to(dev), sc["rotation"].to(dev), sc["opacity"].to(dev)).make_symmetric()
            train_adam = pkg.StoreAdam(train_store, ref_lrs, scale_ceiling=math.log(10.0 * 0.2 + 1e-9), sharded=True)
            train_mode = "fused multimem kernel: reduce-scatter + Adam on the rank's slice + parameter all-gather (csrc/collective.cu)"
        except Exception as e:
            train_store, train_adam = store, adam
            train_mode += f" (symmetric memory unavailable: {type(e).__name__})"

    def step_train(step):
        store = train_store
        store.zero_grad(lazy=True)  # the first backward writes the flat gradient buffer (no memset, no read-modify-write)
        losses = []
        vstreams.begin()
        for v in range(V):
            with vstreams.view(v):
                cur = torch.cuda.current_stream()
                k = train_state["k"]
                slot = k & 1
                if not train_state["primed"]:
                    for sl in range(2):
                        free_ev[sl].record(cur)
                    upload(slot, my_view(step, v))
                    train_state["primed"] = True
                nv, ns = (v + 1, step) if v + 1 < V else (0, step + 1)
                upload(slot ^ 1, my_view(ns, nv))
                cur.wait_event(ready_ev[slot])
                vp = {"world_view_transform": pose_slot[slot], "image_width": W, "image_height": H}
                res = store.render(K, vp, bg_color=bg, render_mode="RGB", sh_degree=deg)
                loss = pkg.l1_ssim_loss(res["render"], gt_slot[slot], 0.2)["loss"] * (1.0 / V)
                loss.backward()
                free_ev[slot].record(cur)
                losses.append(loss.detach())
                train_state["k"] = k + 1
        vstreams.join()
        if world > 1:
            train_adam.step_allreduce(1.0 / world, densify_stats=True)  # fused (sharded) or NCCL + Adam, see train_mode
        else:
            train_adam.step(densify_stats=True)  # K10 rides in the optimiser pass (adam.cu)
        return float(torch.stack(losses).sum().item())

    def timed(fn, steps, warmup, timer=None):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if timer is not None:
            timer.enabled = True
        l0 = pkg._lib.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        torch.cuda.synchronize()
        if timer is not None:
            timer.enabled = False
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, pkg._lib.launch_count - l0

    timer = StageTimer(torch)
    pkg._lib.stage_timer = timer
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev, launches = timed(step_device, args.steps, args.warmup, timer)
    timer_one, ms_one = None, None
    if args.view_streams > 1:
        # the same K steps on ONE stream — what a caller that syncs between views (the unmodified reference trainer,
        # trainer.py:495-497) can get; also gives per-stage durations without the other view's kernels in between
        multi, vstreams = vstreams, pkg.ViewStreams(dev, 1)
        timer_one = StageTimer(torch)
        pkg._lib.stage_timer = timer_one
        ms_one, _ = timed(step_device, args.steps, 3, timer_one)
        vstreams = multi
        pkg._lib.stage_timer = timer
    # multi-GPU diagnostic pass: where does the step go beyond one rank's render?  (a) the collective itself, (b) waiting
    # for the slowest rank.  Events bracket this rank's render and its all-reduce; the collective cannot start before the
    # slowest rank arrives, so the MIN over ranks of the reduce bracket is the collective alone and max-min of the render
    # bracket is the spread.
    multi_gpu = None
    if world > 1:
        probe["on"] = True
        timed(step_device, args.steps, 2)
        probe["on"] = False
        r_ms = sum(a.elapsed_time(b) for a, b in probe["render"][2:]) / max(1, len(probe["render"]) - 2)
        c_ms = sum(a.elapsed_time(b) for a, b in probe["reduce"][2:]) / max(1, len(probe["reduce"]) - 2)
        t = torch.tensor([r_ms, -r_ms, c_ms, -c_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        r_max, r_min, c_max, c_min = float(t[0]), -float(t[1]), float(t[2]), -float(t[3])
        multi_gpu = {"render_ms_max": round(r_max, 4), "render_ms_min": round(r_min, 4), "rank_spread_ms": round(r_max - r_min, 4),
                     "allreduce_ms": round(c_min, 4), "allreduce_bracket_ms_max": round(c_max, 4),
                     "allreduce_bytes": int(sum(p.numel() for p in params_list) * 4),
                     "note": "allreduce_ms = min over ranks of the bracket around the collective (the last rank to arrive waits "
                             "for nobody); rank_spread_ms = max - min over ranks of the per-step render time"}
    ms_e2e, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))
    e2e_mode["async"] = True
    ms_e2e_async, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))
    e2e_mode["async"] = False
    ms_store, _ = timed(step_store, args.steps, max(1, args.warmup // 2))
    ms_train, _ = timed(step_train, args.steps, max(1, args.warmup // 2))
    if rank == 0:
        sampler.stop_flag.set()
        sampler.join(timeout=2)
    pkg._lib.stage_timer = None

    # scene statistics for camera 0 (N_visible, M) and the per-stage roofline
    with torch.no_grad():
        cam = pkg.make_cam(poses_dev[0], K, W, H)
        b = pkg.project_and_bin(leaves["means3D"], leaves["rotations"], leaves["scales"],
                                leaves["opacities"].squeeze(-1).contiguous(), leaves["shs"], deg, cam, W, H)
        M = b["M"]
        n_vis = int((b["radii"] > 0).all(-1).sum().item())
        del b, cam

    # The reference's OWN training step, unmodified (src/training/trainer.py:447-519 as written: getters, autocast,
    # render_with_depth, _stclamp, L1 + five avg_pool2d, three .item() per view, scale regulariser, GradScaler, torch Adam
    # over the six groups), with B200Backend registered behind GaussianRasterizer — one stream, batch_size views, this
    # config's scene.  N=1 only (the reference has no distributed code); needs the staged reference tree.
    ref_trainer = None
    if world == 1 and not args.no_reference_trainer:
        try:
            dh = pkg.dropin  # launcher utilities of the package (no oracle, no tests/ import on this arm)
            if not dh.reference_available():
                raise LookupError("reference tree not staged (baseline/stage_reference.py runs in the build container)")
            start = {k: sc[k] for k in dh.REFERENCE_PARAM_KEYS}
            trainer, model, opt, rast = dh.make_reference_trainer(start, sc["K"], "cuda", batch_size=V)
            trainer.iteration = 10 ** 6  # SH schedule saturated: degree 3, as in every other leg
            assert isinstance(rast.backend, pkg.B200Backend)
            gts = [torch.rand(H, W, 3, generator=gen) for _ in range(2)]
            tviews = dh.tensor_views([cams[i] for i in range(64)], [gts[i & 1] for i in range(64)], sc["K"], W, H, "cuda")

            def step_ref(step):
                trainer._training_step(model, rast, [tviews[my_view(step, v)] for v in range(V)], opt)

            ms_ref, ref_launches = timed(step_ref, args.steps, 3)
            ref_trainer = {"it_per_s": 1e3 / (ms_ref / args.steps), "ms_per_step": ms_ref / args.steps,
                           "Mpix_per_s": V * W * H / (ms_ref / args.steps / 1e3) / 1e6, "views_per_step": V,
                           "b200_kernels_per_step": ref_launches / args.steps,
                           "what": "UNMODIFIED GaussianTrainer._training_step (trainer.py:447-519) + GaussianModel + torch Adam + "
                                   "GradScaler from the staged reference tree, B200Backend registered at "
                                   "gaussian_rasterizer.py:97-118; one stream, host sync every view (.item())"}
        except Exception as e:
            ref_trainer = {"it_per_s": None, "unavailable": f"{type(e).__name__}: {e}"[:200]}

    views_total = args.steps * V
    pix_per_step = world * V * W * H
    value = pix_per_step / (ms_dev / args.steps / 1e3) / 1e6
    e2e_val = pix_per_step / (ms_e2e / args.steps / 1e3) / 1e6

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks, peak_kind = _peaks()
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    tot = timer.totals_ms()
    cnt = timer.counts()
    per_view_ms = {k: tot[k] / max(cnt[k], 1) for k in tot}
    # algorithmic bytes per launch (DESIGN.md "Kernels")
    alg = {
        "project_sh_fwd": n_vis * 280 + (N - n_vis) * 52,
        "bin_emit": M * 12 + N * 16,
        # two-level binning (emit + pair sort + expansion in one C call): the reference formulation's emit (12 B) + sort
        # (24 B compulsory) per intersection, SURVEY §8d
        "bin_coarse": M * 36 + N * 16,
        "sort_pairs": M * 24,
        "tile_ranges": M * 8,
        "tile_count": N * 16,             # radii + means2d of every Gaussian (corner grid, side stream)
        "tile_prep": ((W + 15) // 16 + 1) * ((H + 15) // 16 + 1) * 4 + ((W + 15) // 16) * ((H + 15) // 16) * 4,
        "sort_depth": N * (4 + 4 * 16),   # one histogram read + four onesweep passes over (u32 key, i32 id)
        "bin_scan": N * 8,
        "raster_fwd": M * 44 + W * H * 20,
        "raster_bwd": M * 44 + W * H * 24 + n_vis * 36,
        "project_sh_bwd": n_vis * 508 + (N - n_vis) * 200,
    }
    stages = {}
    for k, ms in per_view_ms.items():
        gbs = alg.get(k, 0) / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        stages[k] = {"ms": round(ms, 4), "alg_MB": round(alg.get(k, 0) / 1e6, 2), "GBps": round(gbs, 1),
                     "frac_hbm": round(gbs / hbm, 4)}
    dom = max(per_view_ms, key=per_view_ms.get) if per_view_ms else None
    roofline = None
    if dom:
        ach = stages[dom]["GBps"]
        traffic = None  # dram bytes per launch of the dominant kernel from the committed ncu capture of this config
        try:
            with open(os.path.join(ROOT, "profiles", f"ncu_traffic_config{args.config}.json")) as f:
                traffic = json.load(f)["bytes"].get(dom)
        except Exception:
            pass
        roofline = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": round(ach / hbm, 4),
                    "traffic": traffic, "binding_resource": "FP32/SFU issue slots (not HBM) for raster_fwd / raster_bwd: 77-79 % issue-active, per-line "
                    "instruction table and lower bound in profiles/ncu_r02_summary.md + profiles/r02/raster_bwd_lines.md" if dom.startswith("raster") else "HBM", "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})",
                    "share_of_step": round(per_view_ms[dom] / sum(per_view_ms.values()), 3), "stages": stages}
        if timer_one is not None:
            t1, c1 = timer_one.totals_ms(), timer_one.counts()
            roofline["stages_note"] = (f"'stages' are CUDA-event brackets inside the timed region, where {args.view_streams} views are in "
                                       "flight on alternating streams: a bracket also holds whatever the other view ran in between. "
                                       "'stages_one_stream_ms' is the same step on ONE stream (separate short pass, not the headline).")
            roofline["stages_one_stream_ms"] = {k: round(t1[k] / max(c1[k], 1), 4) for k in t1}
            # the dominant kernel timed ALONE on its stream (one-stream pass): in the timed region it shares the SMs with the
            # next view's front end (FRONT_PRIORITY), so its launches there last longer than the kernel by itself
            alone_ms = roofline["stages_one_stream_ms"].get(dom)
            if alone_ms:
                ach1 = alg.get(dom, 0) / (alone_ms * 1e-3) / 1e9
                roofline["achieved_alone"] = round(ach1, 1)
                roofline["frac_alone"] = round(ach1 / hbm, 4)

    cpu_baseline = None
    try:
        if world > 1:  # the CPU baseline is measured on rank 0 at N=1 only
            raise LookupError("not measured at N>1 (see the N=1 line)")
        # bounded sample of the same workload on the host cores: one full view, C/OpenMP port of the oracle
        sec, cores, Mc, _ = cpu_port_time(N, W, H, deg, steps=2, warmup=1, views_per_step=V if args.config in ("A", "tiny") else 1)
        cpu_baseline = {"value": (V if args.config in ("A", "tiny") else 1) * W * H / sec / 1e6, "unit": "Mpix/s", "cores": cores, "kind": "port",
                        "sample": f"2 steps x {V if args.config in ('A', 'tiny') else 1} full view(s) fwd+bwd after 1 warm-up step (N={N}, {W}x{H}, M={Mc} at camera 0), oracle/cpu_splat.c OpenMP",
                        "P_pairs_visited": LAST_PAIRS[0] if LAST_PAIRS else None,
                        "P_pairs_composited": LAST_PAIRS[1] if LAST_PAIRS else None}
    except LookupError as e:
        cpu_baseline = {"value": None, "unit": "Mpix/s", "cores": 0, "kind": "port", "sample": str(e)}
    except Exception as e:
        cpu_baseline = {"value": None, "unit": "Mpix/s", "cores": 0, "kind": "port", "sample": f"failed: {e!r}"[:160]}

    line = {
        "impl": "b200",
        "metric": metric_name(args.config), "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.total_views else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.config, V, n_vis, M, world, args.total_views),
        "view_streams": args.view_streams,
        # the same K steps with every view on ONE stream: the figure a caller that host-syncs between views can reach
        "value_one_stream": (pix_per_step / (ms_one / args.steps / 1e3) / 1e6) if ms_one else value,
        "ms_per_step_one_stream": (ms_one / args.steps) if ms_one else ms_dev / args.steps,
        "reference_trainer_it_per_s": (ref_trainer or {}).get("it_per_s"),
        "reference_trainer": ref_trainer,
        "multi_gpu": multi_gpu,
        "gradient_collective": (None if world == 1 else {leg: {
            "off": "five NCCL all-reduces back to back (one per attribute tensor)",
            "nccl": "ONE NCCL all-reduce of a flat buffer the five .grad tensors are views of",
            "sym": "ONE all-reduce by this library's multimem kernel (csrc/collective.cu) over a flat gradient buffer in symmetric memory",
            "arena": "ONE all-reduce by this library's multimem kernel (csrc/collective.cu): autograd's .grad tensors are slices of a "
                     "symmetric buffer handed to the first backward of the step (GradientArena)",
        }.get(m, m) for leg, m in (("value", fg_mode), ("e2e", fg_mode_e2e))}),
        "gradient_arena_hits": ({m: f"{o.hits}/{o.steps} steps reduced by the own collective" for m, o in flat_by_mode.items()
                                if isinstance(o, _Arena)} or None),
        "view_assignment": ("one rank" if world == 1 else ("plain order" if args.no_balance else
                            "balance_views: the step's N*V cameras dealt to the ranks by intersection count M (LPT)")),
        "oracle_pin": gsplat_probe(),
        "e2e": {"value": e2e_val, "unit": "Mpix/s", "h2d_bytes_per_step": V * (W * H * 3 * 4 + 64), "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps},
        # NOT the headline: the same leg with the loss copied to pinned host memory every step without blocking and consumed
        # one step later (a loop that only logs the loss) — shows what the blocking read-back of `e2e` costs
        "e2e_async_readback": {"value": pix_per_step / (ms_e2e_async / args.steps / 1e3) / 1e6, "unit": "Mpix/s",
                               "ms_per_step": ms_e2e_async / args.steps},
        "gpu_launches": launches,
        # second half of BASELINE.json's metric: optimiser-step rate of a loop that renders `views_per_rank` views per
        # rank per step (the reference's batch_size, training_config.py:55); rasterizer fwd+bwd (+ all-reduce) only
        "train_it_per_s": {"device": 1e3 / (ms_dev / args.steps), "e2e": 1e3 / (ms_e2e / args.steps),
                           "views_per_step": world * V},
        "f1_store_path": {"value": pix_per_step / (ms_store / args.steps / 1e3) / 1e6, "unit": "Mpix/s",
                          "ms_per_step": ms_store / args.steps,
                          "what": "same step from GaussianStore (raw parameters, fused activations, in-kernel gradient "
                                  "accumulation into the flat buffer); not the drop-in call, reported beside it"},
        "train_step": {"it_per_s": 1e3 / (ms_train / args.steps), "ms_per_step": ms_train / args.steps,
                       "Mpix_per_s": pix_per_step / (ms_train / args.steps / 1e3) / 1e6, "views_per_step": world * V,
                       "optimizer": train_mode,
                       "what": "full optimiser iteration on the widened path (SURVEY §8 f1+f3+f4): host pose+GT upload, "
                               "GaussianStore.render, fused L1+SSIM loss, backward, all-reduce if N>1, one-kernel Adam, "
                               "densification statistic, loss read back"},
        "clocks": sampler.summary(),
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="A", choices=sorted(CONFIGS))
    ap.add_argument("--views-per-rank", type=int, default=4)
    ap.add_argument("--view-streams", type=int, default=2,
                    help="CUDA streams the views of a step alternate over (ViewStreams); 1 = one stream")
    ap.add_argument("--flat-grads", default="", choices=["", "off", "nccl", "sym", "arena"],
                    help="N>1, drop-in path: 'off' = autograd allocates the five .grad tensors, five NCCL collectives back to back; "
                         "'nccl' = views of one flat buffer, ONE NCCL collective; 'sym' = the flat buffer in symmetric memory, reduced by "
                         "the library's multimem kernel; 'arena' = the first backward of a step writes into slices of a symmetric buffer "
                         "that autograd adopts as .grad (no memset, no extra add), ONE own collective.  Default: 'arena' from 4 ranks "
                         "on, else 'off'")
    ap.add_argument("--no-fused-optimizer", action="store_true",
                    help="N>1, train leg: NCCL all-reduce + Adam on every rank instead of the fused multimem optimiser kernel")
    ap.add_argument("--no-balance", action="store_true",
                    help="N>1: deal the cameras of a step to the ranks in plain order instead of balancing them by M")
    ap.add_argument("--batched-backward", action="store_true", help="value leg: render every view of a step first, then run the backward passes (the reference trainer's order) instead of render+backward per view")
    ap.add_argument("--no-reference-trainer", action="store_true",
                    help="skip the leg that times the reference's unmodified _training_step on B200Backend")
    ap.add_argument("--total-views", type=int, default=0,
                    help="fixed number of views per step split over the ranks (strong scaling, config B: 64); "
                         "default 0 = --views-per-rank views on every rank (weak scaling)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    sys.exit(run_reference(args) if args.impl == "reference" else run_b200(args))


if __name__ == "__main__":
    main()
