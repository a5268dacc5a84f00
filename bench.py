#!/usr/bin/env python
"""bench.py -- path samples/s of the B200 `path` integrator on BASELINE.json's configuration.

Own arm (default):  python bench.py [--gpus N] [--steps K] [--warmup W]
  One "step" = one full render of the workload (every pixel, every sample) through mtsb_render.
  Workload (every N): BASELINE.json configs[2] -- the synthetic 10M-triangle flattened-instance scene, mixed
         diffuse / roughdielectric / roughconductor, constant environment + 4 area lights, 1920x1080, 512 spp,
         maxDepth 10: the configuration north_star states the >=100x target on.  (--workload cfg1_1m / cfg0_cbox /
         cfg4_4k select configs[1] / [0] / [4].)
  N > 1: launched by torchrun, one rank per GPU.  ONE frame of the same size is partitioned by interleaved
         32x32 tiles (tile % N == rank): every rank renders all 512 spp of its tiles ("strong" scaling, the default)
         and one NCCL reduce of the film to rank 0 ends a step.  --scaling weak renders N x 512 spp instead.
  value   = path samples of all ranks / device time (CUDA events inside mtsb_render + the reduce, max over ranks)
  e2e     = same metric through the C ABI with HOST buffers: scene description in host memory ->
            mtsb_scene_create (H2D + BVH build) -> mtsb_render -> mtsb_film_develop (D2H), every step.
Reference arm:      python bench.py --impl reference ...
  The reference's CPU implementation of the same path on the host cores: the kd-tree /
  TriAccel compiled from the reference tree (oracle/_ref) under the restated Li loop (oracle/),
  all host threads, on a bounded sample of the same workload (a subset of 32x32 blocks).
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

METRIC = "path_samples_per_s"
UNIT = "samples/s"
WORKLOADS = {
    # name: (builder name, kwargs, triangles label, closest-hit bytes/ray, shadow bytes/ray)   SURVEY.md 8(d)
    "cfg1_1m": ("scene_1m", dict(width=1920, height=1080, spp=256, max_depth=10), "1M", 720, 708),
    "cfg0_cbox": ("cbox", dict(width=512, height=512, spp=64, max_depth=8), "32", 400, 388),
    "cfg2_10m": ("scene_10m", dict(width=1920, height=1080, spp=512, max_depth=10), "10M", 880, 868),
    # BASELINE configs[4]: the 10M scene at 4K / 1024 spp, meant for 2/4/8 GPUs (tile partition + one film reduce)
    "cfg4_4k": ("scene_10m", dict(width=3840, height=2160, spp=1024, max_depth=10), "10M", 880, 868),
}
SHADE_BYTES = 276      # per shade event


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2_10m", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="N>1: one frame split over the GPUs (strong) or N x spp (weak)")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (testing only; the JSON line says so)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


def make_builder(args):
    from mitsuba_b200 import scenes
    fn, kw, _, _, _ = WORKLOADS[args.workload]
    kw = dict(kw)
    if args.spp > 0:
        kw["spp"] = args.spp
    return getattr(scenes, fn)(**kw), kw


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class CpuReference:
    """The reference's CPU path on the host cores: kd-tree + TriAccel compiled from the reference tree
    (oracle/_ref) under the restated Li loop (oracle/).  Built once; every step() renders a bounded
    sample of the workload (a subset of the 32x32 blocks of the spiral at reduced spp)."""

    def __init__(self, builder, threads=0):
        import oracle                                     # checker / CPU baseline only (never on the product path)
        oracle.build(ref=os.path.isdir("/root/reference/include/mitsuba"))
        # the TIMED build: the reference's release flags at the host's -march (oracle/Makefile FASTFLAGS); parity uses the strict one
        self.flags = oracle.use_timed_build()
        self.use_ref = oracle.ref_available()
        t0 = time.time()
        self.orc = oracle.Oracle(builder, use_ref_kdtree=self.use_ref)
        self.build_s = time.time() - t0
        self.threads = threads
        w, h, _ = self.orc.film_size()
        self.total_blocks = ((w + 31) // 32) * ((h + 31) // 32)
        self.full_spp = builder.integrator.sample_count
        st = self.orc.render(spp=1, seed=1, threads=threads, max_blocks=min(self.total_blocks, 64))
        self.rate = st.samples / max(st.seconds, 1e-6)
        self.kind = "reference" if self.use_ref else "port"

    def step(self, seconds, seed=2):
        want = self.rate * seconds
        blocks = int(min(self.total_blocks, max(min(self.total_blocks, 64), want / (1024.0 * min(self.full_spp, 16)))))
        spp = int(max(1, min(self.full_spp, want / (1024.0 * blocks))))
        st = self.orc.render(spp=spp, seed=seed, threads=self.threads, max_blocks=blocks)
        self.rate = st.samples / max(st.seconds, 1e-6)
        sample = ("%d of %d 32x32 blocks (evenly spaced along the spiral) at %d spp = %d path samples in %.1f s; %s"
                  % (blocks, self.total_blocks, spp, st.samples, st.seconds,
                     "kd-tree + TriAccel compiled from the reference tree under the restated Li loop" if self.use_ref
                     else "CPU restatement (oracle/) with its own BVH"))
        return st.samples / st.seconds, int(st.threads), sample, st

    def record(self, value, cores, sample):
        return {"value": value, "unit": UNIT, "cores": cores, "kind": self.kind, "sample": sample,
                "detail": "kd-tree build/traversal + TriAccel are the reference's own code (compiled from /root/reference); "
                          "Li, BSDFs, emitters, film are the CPU restatement (oracle/): the full mitsuba binary is unbuildable here",
                "flags": self.flags, "kd_build_s": round(self.build_s, 2)}


L2_NOTE = "GPU arm: L2 flushed between steps (160 MB write); per-step working set >> L2"


def workload_config(args, kw, builder, world):
    """The workload-defining keys, identical in both arms (the driver compares them)."""
    strong = args.scaling == "strong"
    total_spp = kw["spp"] if strong else kw["spp"] * world
    tris = int(builder._nt)
    return {"workload": args.workload, **kw, "triangles": tris, "spp_total": total_spp,
            "partition": "tile %% %d == rank, %d spp per rank-tile" % (world, total_spp), "l2": L2_NOTE,
            "spp_override": bool(args.spp)}, total_spp


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    builder, kw = make_builder(args)
    ref = CpuReference(builder)
    per_step = max(2.0, min(20.0, 100.0 / max(1, args.steps + args.warmup)))
    vals, secs = [], []
    cores, sample = 0, ""
    for i in range(args.warmup + args.steps):
        v, cores, sample, st = ref.step(per_step, seed=2 + i)
        if i >= args.warmup:
            vals.append(v)
            secs.append(st.seconds)
    value = float(np.sum([v * s for v, s in zip(vals, secs)]) / np.sum(secs))
    world = max(1, args.gpus)
    config, _ = workload_config(args, kw, builder, world)
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(secs)), "higher_is_better": True,
           "scaling": args.scaling if world > 1 else "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
           "note": "reference CPU path on the host cores; each step is a bounded sample of the workload (see cpu_baseline.sample)",
           "cpu_baseline": ref.record(value, cores, sample),
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def trace_source_hash():
    import hashlib
    h = hashlib.sha256()
    for f in ("traverse.h", "trace.cuh", "bvh.h", "trace.cu"):
        h.update(open(os.path.join(ROOT, "mitsuba_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def host_bytes(builder):
    pos, nrm, uv, idx = builder.arrays()
    return int(pos.nbytes + nrm.nbytes + uv.nbytes + idx.nbytes)


def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    from mitsuba_b200 import lib
    from mitsuba_b200 import dist as mdist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib.init(local)

    builder, kw = make_builder(args)
    # strong (default): ONE frame, tiles interleaved over ranks; weak: N x spp of the same image
    config, total_spp = workload_config(args, kw, builder, world)
    scene = lib.Scene(builder)
    info = scene.info()
    film_t = mdist.film_tensor(scene) if world > 1 else None
    flush = torch.empty(160 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(seed):
        flush.zero_()                                   # L2 flush between timed iterations
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()                              # every rank starts the frame together (outside the timed spans)
            torch.cuda.synchronize()
        st = scene.render(spp=total_spp, seed=seed, tile_index=rank, tile_count=world)
        ms = st.ms_total
        if world > 1:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            mdist.reduce_film(film_t, dst=0)
            e1.record()
            torch.cuda.synchronize()
            ms += e0.elapsed_time(e1)
        return st, ms

    for i in range(args.warmup):
        step(100 + i)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    t_wall0 = time.time()
    dev_ms, stats = 0.0, []
    for i in range(args.steps):
        st, ms = step(200 + i)
        dev_ms += ms
        stats.append(st)
    barrier()
    wall_s = time.time() - t_wall0
    clk = clocks.stop() if rank == 0 else None

    samples_rank = float(sum(s.samples for s in stats))
    agg = torch.tensor([dev_ms, samples_rank, float(sum(s.closest_rays for s in stats)), float(sum(s.shadow_rays for s in stats)),
                        float(sum(s.kernel_launches for s in stats))], dtype=torch.float64, device="cuda")
    mx = agg.clone()
    if world > 1:
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    max_ms = float(mx[0])
    samples_all, rays_all, launches_all = float(agg[1]), float(agg[2] + agg[3]), int(agg[4])
    value = samples_all / (max_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (scene in host memory -> image in host memory) ----
    scene.close()
    desc_pair = builder.desc()          # the scene description as contiguous HOST buffers (the input of the C ABI)
    barrier()
    e2e_t0 = time.time()
    e2e_samples = 0.0
    d2h = 0
    e2e_split = {"create": 0.0, "render": 0.0, "develop": 0.0, "destroy": 0.0}
    for i in range(args.steps):
        t0 = time.time()
        sc = lib.Scene(desc_pair)                                   # H2D of the scene arrays + BVH build
        t1 = time.time()
        st = sc.render(spp=total_spp, seed=300 + i, tile_index=rank, tile_count=world)
        t2 = time.time()
        if world > 1:
            mdist.reduce_film(mdist.film_tensor(sc), dst=0)
            torch.cuda.synchronize()
        if rank == 0:
            img = sc.film_develop()                                 # D2H of the developed image
            d2h = int(img.nbytes)
        t3 = time.time()
        e2e_samples += st.samples
        sc.close()
        t4 = time.time()
        for k, dt in zip(("create", "render", "develop", "destroy"), (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            e2e_split[k] += dt * 1e3 / args.steps
    barrier()
    e2e_s = time.time() - e2e_t0
    e2e_all = torch.tensor([e2e_samples], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_all, op=dist.ReduceOp.SUM)
    e2e_value = float(e2e_all[0]) / e2e_s

    if rank == 0:
        _, _, tris, b_c, b_s = WORKLOADS[args.workload]
        s0 = stats[-1]
        # roofline of the dominant kernel (closest-hit traversal): algorithmic bytes / launch time
        closest = float(sum(s.closest_rays for s in stats)); ms_closest = float(sum(s.ms_trace_closest for s in stats))
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = closest * b_c / (ms_closest * 1e-3) / 1e9 if ms_closest > 0 else 0.0
        # measured DRAM traffic of the same kernel: only when an ncu --set full capture of THIS kernel source and workload
        # is committed (profiles/r02_traffic.json records the hash of the traversal sources it was taken on)
        traffic, traffic_note = None, "no ncu capture of this kernel build / workload committed"
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
            ent = tj.get(args.workload)
            if ent and ent.get("trace_src_sha16") == trace_source_hash():
                n_launch = max(1, int(sum(s.closest_launches for s in stats)))
                traffic = ent["dram_bytes_per_ray_closest"] * closest / n_launch
                traffic_note = ("DRAM bytes per k_trace<closest> launch = measured bytes/ray (%s) x mean rays per launch; algorithmic bytes/ray = %d"
                                % (ent.get("source", "ncu --set full"), b_c))
        except Exception:
            pass
        n_closest_launches = max(1, int(sum(s.closest_launches for s in stats)))
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": max_ms / args.steps, "higher_is_better": True, "scaling": args.scaling if world > 1 else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "detail": {"bvh_nodes": int(info.n_bvh_nodes), "bvh_build_ms": round(float(info.build_ms), 2)},
            "mrays_per_s": rays_all / (max_ms * 1e-3) / 1e6,
            "rays_per_sample": rays_all / samples_all,
            "wall_s_timed_region": wall_s,
            "stage_ms_last_step": {"raygen": s0.ms_raygen, "trace_closest": s0.ms_trace_closest, "shade": s0.ms_shade,
                                   "trace_any": s0.ms_trace_any, "splat": s0.ms_splat, "total": s0.ms_total, "passes": s0.n_passes},
            # rays in flight per bounce of the last step (this rank's tiles): closest-hit rays of bounce b (b = 0: camera rays), shadow rays made at bounce b
            "rays_in_flight_per_bounce": {"closest": [int(x) for x in s0.closest_rays_bounce if x], "shadow": [int(x) for x in s0.shadow_rays_bounce if x]},
            "roofline": {"bound": "hbm", "kernel": "k_trace<closest>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)",
                         "bytes_per_ray": b_c, "launches": n_closest_launches, "avg_launch_ms": ms_closest / n_closest_launches,
                         "rays_per_launch": closest / n_closest_launches},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": host_bytes(builder), "d2h_bytes_per_step": d2h,
                    "includes": "mtsb_scene_create (H2D + GPU BVH build) + mtsb_render + film reduce + mtsb_film_develop (D2H) per step",
                    "wall_ms_per_step": {k: round(v, 2) for k, v in e2e_split.items()}},
            "gpu_launches": launches_all, "clocks": clk,
        }
        if not args.no_cpu_baseline and world == 1:
            ref = CpuReference(builder)
            v, cores, sample, _ = ref.step(args.cpu_seconds)
            out["cpu_baseline"] = ref.record(v, cores, sample)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)
