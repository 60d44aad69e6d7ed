#!/usr/bin/env python
"""Benchmark of the batched anomaly generator (BASELINE.json metric: synthetic-anomaly images/s and
HBM GB/s, vs the reference CPU path).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 1|2|3|4|5]

A step = one pass of the hot path over one batch of synthetic inputs (config 4: ssad_obj_mask + ssad_generate).
Headline workload at N=1 = BASELINE.json configs[3] -- the largest single-GPU configuration: 1024 images of
1024x1024, non-fixed-pose class, per-image obj_mask + generation (BASELINE.json's metric names no config).  The same
JSON line carries `per_config`: value / ms_per_step / roofline.frac / e2e / verify for configs 1, 2, 3 (and the
config-5 trainer feed), each measured the same way in the same run.  Multi-GPU: the batch shards by image, every rank
runs the same per-GPU batch (weak scaling), no collective on the data path.

Prints ONE JSON line (rank 0).  `value` = images/s with inputs resident in HBM; `e2e` = the same through host buffers
(fresh parameter records drawn by the host sampler every step, pinned H2D of images + records, D2H of the uint8
result + labels, all inside the timed region); `roofline` for the step's kernels; `verify` = images of the TIMED
output buffers compared bit for bit with the CPU oracle after the timed region; `cpu_baseline` = the CPU oracle (the
reference's algorithm on the installed Pillow/torchvision) timed on this box's host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC, UNIT = "synthetic_anomaly_images_per_sec", "images/s"
DEFAULT_BATCH = {1: 96, 2: 512, 3: 1024, 4: 1024, 5: 512}
KERNELS = {
    "generate": "ssad_generate: plan, lists, pass1, canvas, finish, pass2 (_v7_kernel), timed as one step",
    "obj_mask": "ssad_obj_mask (canny32, canny_nms, canny_resolve, mask_planes) + ssad_generate (6 _v7_kernel), timed as one step",
}


# ------------------------------------------------------------------ workloads
def make_workload(config: int, batch: int, seed: int = 0):
    """Returns (cfg, images u8 [B,S,S,3], mask bool [S,S] | None, donors | None, params, name, AB)."""
    import ssad_b200
    from ssad_b200 import GeneratorConfig, sample_params, synth
    uniq = 16   # distinct synthetic images; the batch cycles through them (MVTec classes have 60-391)
    if config == 2:
        S, ps = 320, 256
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S), patch_localization=True, patch_size=ps,
                              patch_area_ratio=(0.1, 0.2), scar_area_ratio=(0.01, 0.02))
        base = synth.object_images(uniq, S, S, 1234 + config)
        mask = synth.disc_mask(S, S, 0.5)
        donors = None
        name = "patch-level 3-class, 256x256 windows of 320x320 images, area_ratio [0.1,0.2]/[0.01,0.02]"
        ab = 2 * 3 * ps * ps + ps * ps // 8
    elif config in (1, 5):
        S = 256
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        base = synth.object_images(uniq, S, S, 1234 + config)
        mask = synth.disc_mask(S, S, 0.4)
        donors = None
        name = "image-level 3-class, 256x256 fixed-pose object, RandomAffine pre-warp"
        ab = 2 * 3 * S * S
        if config == 5:
            name += ", bf16 NCHW output handed to the ResNet-18 trainer on the same stream"
            ab = 3 * S * S + 6 * S * S
    elif config == 3:
        S = 256
        cfg = GeneratorConfig(subject="carpet", imsize=(S, S), n_donors=15)
        base = synth.texture_images(uniq, S, S, 1234 + config, tint=(120, 110, 100))
        donors = np.concatenate([synth.texture_images(1, S, S, 5000 + k, tint=((37 * k) % 200 + 30, (91 * k) % 200 + 30, (53 * k) % 200 + 30))
                                 for k in range(15)])
        mask = None
        name = "texture class, 256x256, cross-class donor bank (15), all-positive mask, RandomAffine pre-warp"
        ab = 2 * 3 * S * S
    elif config == 4:
        S = 1024
        cfg = GeneratorConfig(subject="screw", imsize=(S, S))
        uniq = 8
        base = synth.screw_images(uniq, S, S, 1234 + config)
        mask = None                          # per image, rebuilt on the device every step (ssad_obj_mask)
        donors = None
        name = "non-fixed-pose object (screw-like), 1024x1024, per-image obj_mask recompute + generation"
        ab = 2 * 3 * S * S + 2 * S * S
    else:
        raise SystemExit(f"config {config} is not a bench line")
    images = base[np.arange(batch) % uniq]
    params = sample_params(cfg, batch, np.random.default_rng(5678 + config + seed))
    return cfg, images, mask, donors, params, name, ab


def config_block(config: int, name: str, batch: int, world: int, out: str, inflight: int, graph: bool) -> dict:
    """The `config` object of a JSON line -- the same keys for our arm and for the reference arm."""
    return {"workload": name, "baseline_config": config, "batch_per_gpu": batch, "global_batch": batch * world, "output": out,
            "l2": "two alternating resident input/output sets; one step moves more than the 126 MB L2",
            "inflight": inflight, "cuda_graph": bool(graph), "parallelism": f"image-sharded x{world}, no collective"}


# ------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.005):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.sm_max = [], set(), None
        self._halt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def _phys_gpu(index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis and all(t.strip().isdigit() for t in vis.split(",")):
        return int(vis.split(",")[index])
    return index


# ------------------------------------------------------------------ CPU oracle: timing and verification
_W = {}


def _cpu_batch(config, batch):
    """Images each CPU worker holds: the whole batch, or a slice of it where the batch is gigabytes (config 4)."""
    return min(batch, 16) if config == 4 else batch


def _cpu_init(config, batch):
    """Worker start-up: one torch / BLAS thread per process (as DataLoader workers run),
    the workload, and one untimed item so that imports (sklearn, torchvision) are outside the timed region."""
    import torch
    torch.set_num_threads(1)
    try:
        from threadpoolctl import threadpool_limits
        _W["tp"] = threadpool_limits(1)
    except Exception:
        pass
    import scenarios as S
    cfg, images, mask, donors, params, _, _ = make_workload(config, _cpu_batch(config, batch))
    _W["sc"] = S.Scenario("bench", cfg, images, mask, donors, params)
    _cpu_run(0)


def _oracle_item(sc, i):
    """The CPU oracle for item i of a scenario -> (x u8 [h,w,3], y).  Non-fixed classes rebuild obj_mask per image
    (dataset_generator.py:27-39 as restated in oracle/objmask.py), exactly what the device step does."""
    import scenarios as S
    from PIL import Image
    from oracle import generator as og
    if sc.cfg.is_non_fixed and sc.mask is None:
        from oracle import objmask
        x, y, _ = og.generate_item(Image.fromarray(sc.images[i]), sc.params.image_draws(i), sc.cfg, None, None, objmask.obj_mask)
        return np.array(x), int(y)
    x, y, _ = S.run_oracle(sc, i)
    return x, int(y)


def _cpu_run(i):
    from PIL import Image
    from oracle import generator as og
    x, y = _oracle_item(_W["sc"], i)
    og.to_tensor_normalize(Image.fromarray(x))      # datasets.py:392-393, part of the path
    return y


def cpu_oracle_images_per_s(config: int, batch: int, sample: int, procs: int, target_s: float = 0.0):
    """images/s of the CPU oracle over `sample` images of the workload on `procs` processes (each worker is
    warmed by its initializer).  With target_s > 0 the sample is re-sized from a pilot run so that the timed
    region lasts about target_s seconds.  Returns (images/s, seconds, images)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(procs, initializer=_cpu_init, initargs=(config, batch)) as pool:
        def run(n):
            t0 = time.perf_counter()
            pool.map(_cpu_run, [i % _cpu_batch(config, batch) for i in range(n)], chunksize=max(1, n // (procs * 8)))
            return time.perf_counter() - t0
        run(min(sample, 2 * procs))          # untimed: first-call costs of every worker (page faults, lazy imports)
        dt = run(sample)
        if target_s > 0:
            sample = int(max(sample, min(50000, target_s * sample / dt)))
            dt = run(sample)
    return sample / dt, dt, sample


def _verify_init(sc):
    import torch
    torch.set_num_threads(1)
    _W["vsc"] = sc


def _verify_run(i):
    return _oracle_item(_W["vsc"], i)


def verify_against_oracle(cfg, images, mask, donors, params, got_x, got_y, idx, norm_dtype=None, procs=1):
    """Compares device results (numpy: got_x [n,h,w,3] u8 or [n,3,h,w] normalised, got_y [n]) for the items `idx` of the
    workload with the CPU oracle on the same records.  Returns the number of mismatching images."""
    import scenarios as S
    sc = S.Scenario("verify", cfg, images, mask, donors, params)
    if procs > 1 and len(idx) > 1:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(min(procs, len(idx)), initializer=_verify_init, initargs=(sc,)) as pool:
            res = pool.map(_verify_run, [int(i) for i in idx])
    else:
        res = [_oracle_item(sc, int(i)) for i in idx]
    bad = 0
    for k, (x, y) in enumerate(res):
        ok = int(got_y[k]) == y
        if norm_dtype is None:
            ok = ok and np.array_equal(got_x[k], x)
        else:
            import torch
            from PIL import Image
            from oracle import generator as og
            want = og.to_tensor_normalize(Image.fromarray(x)).to(norm_dtype).float().numpy()
            ok = ok and np.array_equal(got_x[k], want)
        bad += 0 if ok else 1
    return bad


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


_FULL_AFFINITY = []          # the process's CPU set before _bind_near_gpu narrowed it (restored for the CPU legs)


def _bind_near_gpu(index: int):
    """Bind this process to the CPUs NVML reports as closest to GPU `index` (one process per GPU: the pinned host
    buffers of the end-to-end leg are then first-touched on that NUMA node).  Returns the number of CPUs bound, or None."""
    if os.environ.get("SSAD_BENCH_NO_AFFINITY"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(_phys_gpu(index))
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        full = os.sched_getaffinity(0)
        cpus &= full
        if cpus:
            _FULL_AFFINITY.append(full)
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def _restore_affinity():
    if _FULL_AFFINITY:
        os.sched_setaffinity(0, _FULL_AFFINITY[0])


class ParamPrefetcher:
    """Draws the parameter records on a background thread while the current batch runs (the host sampler is part of the
    end-to-end path: SURVEY.md section 8d/8e).  Small batches are drawn several at a time (one call of the sampler costs
    ~60 NumPy calls whatever the batch) and handed out as slices; every step still gets records nobody has used."""
    POOL = 2048

    def __init__(self, cfg, batch, seed):
        from ssad_b200 import sample_params
        self.cfg, self.batch, self.rng, self.sample = cfg, batch, np.random.default_rng(seed), sample_params
        self.per_pool = max(1, self.POOL // batch)
        self.busy_s, self.n = 0.0, 0
        self._next, self._thread = None, None
        self._cur, self._used = None, 0
        self._start()

    def _work(self):
        t0 = time.perf_counter()
        self._next = self.sample(self.cfg, self.batch * self.per_pool, self.rng)
        self.busy_s += time.perf_counter() - t0
        self.n += self.batch * self.per_pool

    def _start(self):
        self._thread = threading.Thread(target=self._work, daemon=True)
        self._thread.start()

    def get(self):
        if self._cur is None or self._used == self.per_pool:
            self._thread.join()
            self._cur, self._used = self._next, 0
            self._start()
        p = self._cur if self.per_pool == 1 else self._cur.slice(self._used * self.batch, (self._used + 1) * self.batch)
        self._used += 1
        return p

    def close(self):
        self._thread.join()


# ------------------------------------------------------------------ one configuration on this rank's GPU
def run_config(config: int, args, rt, steps: int, e2e_steps: int, verify_images: int, batch=None, out=None):
    """Times one BASELINE configuration.  rt = dict(rank, world, local_rank, dev, dist).  Returns the measured fields."""
    import torch
    from ssad_b200.generator import BatchGenerator, obj_mask_batch
    dev, world, rank, dist = rt["dev"], rt["world"], rt["rank"], rt["dist"]
    batch = batch or DEFAULT_BATCH[config]
    out = out or args.out
    cfg, images, mask, donors, params, name, ab = make_workload(config, batch, seed=rank)
    gen = BatchGenerator(cfg, dev, mask=mask, donors=donors)
    W, H = cfg.work_size
    norm_dtype = {"u8": None, "f32": torch.float32, "bf16": torch.bfloat16}[out]
    want_u8 = out == "u8"
    if out == "f32":
        ab = ab - 3 * H * W + 12 * H * W
    elif out == "bf16":
        ab = ab - 3 * H * W + 6 * H * W
    # two resident input sets, alternated, so consecutive steps never touch the same lines
    # (one step already moves > L2: e.g. config 2 = 157 MB in + 100 MB out vs 126 MB L2)
    # batches in flight: two; four for batches under 256 images (the reference's B = 96: one such batch is six short
    # kernels in a row on a fraction of the SMs, a latency chain that more batches in flight hide)
    inflight = args.inflight if args.inflight else (4 if batch < 256 else 2)
    inflight = max(1, min(4, inflight))
    nset = max(2, inflight)                  # input / output / workspace sets (one per batch in flight, two at least)
    srcs = [torch.from_numpy(images).to(dev)]
    srcs.append(torch.roll(srcs[0], 1, dims=0).contiguous())
    srcs += [srcs[k & 1] for k in range(2, nset)]         # (sets beyond the second read the same two inputs)
    per_image_masks = cfg.is_non_fixed
    mask_bufs, mask_ws = [None] * 4, None
    if per_image_masks:
        mask_bufs = [obj_mask_batch(srcs[j]) for j in range(nset)]
        mask_ws = [torch.empty((int(gen.lib.ssad_obj_mask_workspace(batch, images.shape[1], images.shape[2])) + 16,),
                               dtype=torch.uint8, device=dev) for _ in range(nset)]
    pdev = gen.upload_params(params)
    outs_u8 = [torch.empty((batch, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(nset)] if want_u8 else [None] * nset
    outs_n = [torch.empty((batch, 3, H, W), dtype=norm_dtype, device=dev) for _ in range(nset)] if norm_dtype else [None] * nset
    y_outs = [torch.empty((batch,), dtype=torch.int64, device=dev) for _ in range(nset)]

    plans = [gen.plan(srcs[j], params, masks=mask_bufs[j], want_u8=want_u8, norm_dtype=norm_dtype, params_dev=pdev,
                      out_u8=outs_u8[j], out_norm=outs_n[j], y_out=y_outs[j]) for j in range(nset)]
    om_launches = int(gen.lib.ssad_obj_mask_launches()) if per_image_masks else 0
    launches_per_step = int(gen.lib.ssad_generate_launches(plans[0].args[0], int(bool(args.graph)))) + om_launches
    if args.graph:
        for pl in plans:
            gen.capture(pl)                      # one cudaGraphLaunch per step instead of a dozen kernel launches

    def step(i):
        if per_image_masks:                  # obj_mask of every image of the batch, then the generator (datasets.py:232-233)
            obj_mask_batch(srcs[i % nset], out=mask_bufs[i % nset], workspace=mask_ws[i % nset])
        gen.run(plans[i % nset])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # Batches in flight: consecutive steps alternate between `--inflight` streams (each step = the complete hot path over
    # one batch, with its own input set, outputs and workspace), the way a loader keeps a second batch in preparation
    # while the consumer takes the first.  --inflight 1 issues every step on one stream (also timed, `serial` below).
    cur = torch.cuda.current_stream(dev)
    lanes = [torch.cuda.Stream(dev) for _ in range(inflight)]

    def timed(nsteps, use_lanes):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(cur)
        if use_lanes and len(lanes) > 1:
            for ln in lanes:
                ln.wait_stream(cur)
            for i in range(nsteps):
                with torch.cuda.stream(lanes[i % len(lanes)]):
                    step(i)
            for ln in lanes:
                cur.wait_stream(ln)
        else:
            for i in range(nsteps):
                step(i)
        e1.record(cur)
        barrier()
        return e0.elapsed_time(e1)

    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    barrier()
    if getattr(args, "profile_steps", 0) > 0:      # the steps a profiler looks at (plain order on one stream), nothing else
        torch.cuda.profiler.start()
        for i in range(args.profile_steps):
            step(i)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return {"profiled_steps": args.profile_steps, "launches_per_step": launches_per_step}
    timed(warm, True)                        # warm the lanes too
    ms_serial = timed(steps, False)
    sampler = ClockSampler(_phys_gpu(rt["local_rank"]))
    sampler.start()
    ms = timed(steps, True)
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / steps
    res = {"config": config, "name": name, "batch": batch, "out": out, "ab": ab, "steps": steps, "warmup": warm,
           "ms_per_step": ms_per_step, "value": world * batch / (ms_per_step * 1e-3), "clocks": clocks,
           "serial_ms_per_step": ms_serial / steps, "launches_per_step": launches_per_step, "inflight": len(lanes)}
    if per_image_masks:                      # mask builder and generator on their own
        parts = {}
        for nm, fn in (("obj_mask_ms", lambda i: obj_mask_batch(srcs[i & 1], out=mask_bufs[i & 1], workspace=mask_ws[i & 1])),
                       ("generate_ms", lambda i: gen.run(plans[i & 1]))):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            barrier()
            parts[nm] = e0.elapsed_time(e1) / steps
        res["parts"] = parts

    # ---- verify: images of the TIMED output buffers (both alternating sets, as the graph replays left them) against
    # the CPU oracle on the same records -- bit-exact, after the timed region
    if verify_images > 0 and rank == 0:
        n_each = max(1, verify_images // 2)
        bad, checked = 0, 0
        t0 = time.perf_counter()
        _restore_affinity()
        for j in range(2):
            idx = np.unique(np.linspace(0, batch - 1, n_each).astype(np.int64))
            gx = (outs_u8[j] if want_u8 else outs_n[j].float())[torch.as_tensor(idx, device=dev)].cpu().numpy()
            gy = y_outs[j][torch.as_tensor(idx, device=dev)].cpu().numpy()
            imgs_j = images if j == 0 else np.roll(images, 1, axis=0)
            bad += verify_against_oracle(cfg, imgs_j, mask, donors, params, gx, gy, idx, norm_dtype,
                                         procs=host_cores() if config == 4 else 1)
            checked += len(idx)
        res["verify"] = {"images": checked, "mismatch": bad, "seconds": round(time.perf_counter() - t0, 2),
                         "what": "timed output buffers of both alternating sets vs oracle/generator.py on the same records, bit-exact"}

    # ---- e2e: host buffers in, host buffers out, fresh records from the host sampler, all copies inside the timed region
    if e2e_steps > 0:
        h_src = [torch.from_numpy(images).pin_memory()]
        h_src.append(h_src[0] if images.nbytes > (1 << 30) else torch.from_numpy(np.roll(images, 1, axis=0).copy()).pin_memory())
        h_out = torch.empty((batch, H, W, 3), dtype=torch.uint8).pin_memory()
        h_y = torch.empty((batch,), dtype=torch.int64).pin_memory()
        rec_bytes = params.records.nbytes
        masks_fn = (lambda t: obj_mask_batch(t)) if per_image_masks else None
        deep = h_out.numel() <= (1 << 30)        # (config 4's 3 GB result keeps one host set: every call waits)
        h_outs = [h_out, torch.empty_like(h_out).pin_memory() if deep else h_out]
        h_ys = [h_y, torch.empty_like(h_y).pin_memory() if deep else h_y]
        pending = [None, None]
        host_s = [0.0]                           # host time spent inside the call (enqueueing), summed
        pre = ParamPrefetcher(cfg, batch, 99 + rank)

        def e2e_step(i):
            # the public host-buffer call: pieces of the batch flow through H2D copy -> ssad_generate -> D2H copy streams.
            # Two host result sets alternate; a step first takes delivery of the batch that used its set two steps ago
            # (the consumer), then issues its own (wait=False), so one batch is always in flight behind the one delivered.
            j = i & 1
            if pending[j] is not None:
                pending[j].synchronize()
            p = pre.get()                        # this step's records: drawn while the previous step was being issued
            t_in = time.perf_counter()
            pending[j] = gen.generate_from_host(h_src[j], p, h_outs[j], h_ys[j], chunks=args.e2e_chunks,
                                                masks_fn=masks_fn, wait=not deep, upload=args.e2e_upload)
            host_s[0] += time.perf_counter() - t_in

        for i in range(3):
            e2e_step(i)
        barrier()
        import gc
        gc.collect()                             # (buffers of the previous configuration are released here, not in the timed region)
        # the host side of this leg (Python issuing, the sampler thread, pinned-memory traffic of a shared box) makes
        # single runs of a few dozen milliseconds jumpy: the steps are timed in three stretches, the MEDIAN stretch is
        # the value and all three are listed (one stretch for config 4: its four steps take 0.3 s)
        reps = 3 if deep else 1
        per_rep = max(1, e2e_steps // reps)
        host_s[0] = 0.0
        rep_dt = []
        for _ in range(reps):
            t0 = time.perf_counter()
            for i in range(per_rep):
                e2e_step(i)
            barrier()
            dt_ = time.perf_counter() - t0
            if world > 1:
                t = torch.tensor([dt_], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt_ = float(t.item())
            rep_dt.append(dt_)
        pre.close()
        e2e_steps_run = per_rep * reps
        e2e_dt = sorted(rep_dt)[len(rep_dt) // 2]
        res["e2e"] = {"value": world * batch * per_rep / e2e_dt, "unit": UNIT,
                      "stretches": [world * batch * per_rep / d_ for d_ in rep_dt],
                      "h2d_bytes_per_step": int(images.nbytes + rec_bytes),
                      "d2h_bytes_per_step": int(h_out.numel() + h_y.numel() * 8), "steps": e2e_steps_run,
                      "in_flight": 2 if deep else 1, "cpus_bound": rt.get("numa"),
                      "host_ms_per_call": 1e3 * host_s[0] / max(e2e_steps_run, 1),
                      "sampler_included": True,
                      "sampler_records_per_s": (pre.n / pre.busy_s) if pre.busy_s > 0 else None}
        # ---- the same with the source set RESIDENT on the device (the DataModule's deployment, datasets.py:476-508: the
        # decoded set is loaded once): per step only the records go up and the results come down; per-image masks are
        # per SOURCE image and built once (obj_mask depends on the original only)
        if not args.no_resident:
            uniq_n = min(batch, 8 if config == 4 else 16)     # (make_workload cycles the batch through that many images)
            uniq_dev = torch.from_numpy(np.ascontiguousarray(images[:uniq_n])).to(dev)
            idx_sets = [np.arange(batch) % uniq_n, (np.arange(batch) + 1) % uniq_n]
            src_masks = obj_mask_batch(uniq_dev) if per_image_masks else None
            pre2 = ParamPrefetcher(cfg, batch, 199 + rank)
            pending2 = [None, None]

            def res_step(i):
                j = i & 1
                if pending2[j] is not None:
                    pending2[j].synchronize()
                pending2[j] = gen.generate_resident_to_host(uniq_dev, idx_sets[j], pre2.get(), h_outs[j], h_ys[j], masks=src_masks,
                                                            chunks=args.e2e_chunks, wait=not deep)
            for i in range(3):
                res_step(i)
            barrier()
            t0 = time.perf_counter()
            for i in range(e2e_steps):
                res_step(i)
            barrier()
            dt2 = time.perf_counter() - t0
            pre2.close()
            if world > 1:
                t = torch.tensor([dt2], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt2 = float(t.item())
            res["e2e_resident"] = {"value": world * batch * e2e_steps / dt2, "unit": UNIT, "h2d_bytes_per_step": int(rec_bytes),
                                   "d2h_bytes_per_step": int(h_out.numel() + h_y.numel() * 8), "steps": e2e_steps,
                                   "resident_images": int(uniq_n), "sampler_included": True,
                                   "what": "source set resident on the device (decoded once, as the DataModule keeps it), records up, "
                                           "results down" + ("; one obj_mask per SOURCE image, built once" if per_image_masks else "")}
            del uniq_dev, src_masks
        del h_src, h_out, h_outs, h_y, h_ys
    del srcs, outs_u8, outs_n, plans, gen, mask_bufs, mask_ws, pdev
    torch.cuda.empty_cache()
    return res


def datamodule_rate(rt, items: int = 1000, batch: int = 96, epochs: int = 3) -> dict:
    """VERDICT item 4 / datasets.py:501-508: what replaces ``DataLoader(PretextTaskDataset, batch_size=96, shuffle=True,
    num_workers=8)`` -- one epoch of ``BatchLoader`` over a device-resident ``PretextTaskDataset`` of 1000 items (40
    distinct synthetic images, duplicated as ``duplicate_filenames`` does), records drawn by the host sampler on the
    loader's background thread, float32 NCHW tensors + labels (+ the originals, as the reference returns them)."""
    import torch
    from ssad_b200 import synth
    from ssad_b200.datasets import BatchLoader, PretextTaskDataset
    dev = rt["dev"]
    S = 256
    base = synth.object_images(40, S, S, 4242)
    ds = PretextTaskDataset("bottle", imsize=(S, S), images=base, fixed_segmentation=synth.disc_mask(S, S, 0.4), device=dev,
                            seed=7 + rt["rank"], dataset_dir="/nonexistent/")
    ds.images_filenames = np.arange(items)                 # 1000 items over the 40 resident images
    ds.index = np.arange(items) % len(base)
    loader = BatchLoader(ds, batch, shuffle=True, drop_last=True)
    n = 0
    for b in loader:                                       # warm-up epoch (workspace, contexts)
        n += 1
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    seen = 0
    for _ in range(epochs):
        for b in loader:
            seen += int(b[1].shape[0])
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    return {"value": seen / dt, "unit": UNIT, "items": items, "batch": batch, "batches_per_epoch": n, "epochs": epochs,
            "what": "PretextTaskDataset (device-resident) + BatchLoader, shuffle, drop_last, host sampler included, float32 NCHW + labels"}


def roofline_of(res: dict, peaks: dict, traffic: dict) -> dict:
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = res["batch"] * res["ab"] / (res["ms_per_step"] * 1e-3) / 1e9      # per GPU: one step = one batch
    return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic.get(f"config{res['config']}_{res['out']}"),
            "kernel": KERNELS["obj_mask" if res["config"] == 4 else "generate"],
            "algorithmic_bytes_per_image": res["ab"], "algorithmic_bytes_per_step": res["batch"] * res["ab"],
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s"}


# ------------------------------------------------------------------ reference arm: the CPU implementation of the path
def run_reference(args, rank: int):
    """`--impl reference`: the reference's algorithm for the path (oracle/generator.py = the reference's statements on the
    installed Pillow / torchvision; the reference is pure Python with no installable package) on every
    host core, each step a bounded sample of the same workload.  Rank 0 only."""
    if rank != 0:
        return
    config = args.config
    batch = args.batch or DEFAULT_BATCH[config]
    cores = host_cores()
    _, _, _, _, _, name, _ = make_workload(config, 8)
    # bound the run: ~90 images/s/core expected (1.5 with the 1024x1024 obj_mask); keep (steps + warmup) * sample under ~2 min
    per_core = 1.5 if config == 4 else 60.0
    steps, warm = (args.steps or 20), args.warmup
    sample = int(max(8, min(batch, 120.0 * per_core * cores / max(1, steps + warm))))
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_cpu_init, initargs=(config, batch)) as pool:
        idx = [i % _cpu_batch(config, batch) for i in range(sample)]
        for _ in range(warm):
            pool.map(_cpu_run, idx, chunksize=max(1, sample // (cores * 4)))
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_cpu_run, idx, chunksize=max(1, sample // (cores * 4)))
        dt = time.perf_counter() - t0
    v = sample * steps / dt
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_block(config, name, batch, world, args.out, max(1, min(4, args.inflight or (4 if batch < 256 else 2))), bool(args.graph)),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{sample} images/step of the workload, {cores} processes x 1 thread; the reference is "
                                       "pure Python and cannot travel to this box, so its algorithm runs as oracle/generator.py "
                                       "on the installed Pillow/torchvision"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------ main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps of the headline config (default 30 for config 4, else 300)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=4, help="headline configuration (BASELINE.json configs[n-1]); default 4")
    ap.add_argument("--batch", type=int, default=None, help="per-GPU batch of the headline config (default: the config's)")
    ap.add_argument("--out", default="u8", choices=["u8", "f32", "bf16"], help="output variant timed for `value`")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--per-config", type=int, default=None, help="1: also measure the other BASELINE configs into `per_config` "
                    "(default: on for a 1-GPU run of the default headline)")
    ap.add_argument("--verify", type=int, default=32, help="images of the timed buffers compared with the CPU oracle (0: off)")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--e2e-upload", default="auto", choices=["auto", "full", "windows"],
                    help="host-buffer call: whole images up (auto = full), or only the windows a patch-mode call reads")
    ap.add_argument("--no-resident", action="store_true", help="skip the e2e_resident leg")
    ap.add_argument("--e2e-chunks", type=int, default=None, help="pieces of the batch in flight in the host-buffer call")
    ap.add_argument("--inflight", type=int, default=None, help="batches in flight: steps rotate over this many streams and buffer sets "
                    "(1..4; default 2, 4 for batches under 256 images)")
    ap.add_argument("--graph", type=int, default=1, help="1: replay each step's ssad_generate call from a CUDA graph")
    ap.add_argument("--profile-steps", type=int, default=0, help="run this many steps of --config between cudaProfilerStart/Stop "
                    "and exit (for `ncu --profile-from-start off`: tools/traffic.py)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = _bind_near_gpu(local_rank)            # host threads + pinned buffers on the GPU's own NUMA node (e2e copies)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rt = {"rank": rank, "world": world, "local_rank": local_rank, "dev": dev, "dist": dist, "numa": numa}

    if args.config == 5:
        from ssad_b200 import trainer_feed
        line = trainer_feed.bench_line(args, rt, make_workload, config_block, ClockSampler(_phys_gpu(local_rank)))
        if rank == 0:
            print(json.dumps(line))
        if world > 1:
            dist.destroy_process_group()
        return

    steps = args.steps or (30 if args.config == 4 else 300)
    e2e_steps = args.e2e_steps if args.e2e_steps is not None else (4 if args.config == 4 else 20)
    head = run_config(args.config, args, rt, steps, e2e_steps, args.verify, batch=args.batch)
    if "profiled_steps" in head:
        print(json.dumps(head))
        return
    per_config = {}
    want_pc = args.per_config if args.per_config is not None else int(world == 1 and args.config == 4 and args.batch is None)
    if want_pc:
        for c in (1, 2, 3):
            if c == args.config:
                continue
            r = run_config(c, args, rt, max(steps, 200), 30, args.verify)
            per_config[str(c)] = r
        try:
            per_config["datamodule"] = datamodule_rate(rt)
        except Exception as e:
            per_config["datamodule"] = {"error": f"{type(e).__name__}: {e}"}
        try:
            from ssad_b200 import trainer_feed
            per_config["5"] = trainer_feed.measure(args, rt, make_workload, steps=20, warmup=5)
        except Exception as e:                    # the trainer feed needs torchvision's ResNet-18; report, do not hide
            per_config["5"] = {"error": f"{type(e).__name__}: {e}"}
    _restore_affinity()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks, traffic = {}, {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": head["steps"], "warmup": head["warmup"],
        "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": config_block(head["config"], head["name"], head["batch"], world, head["out"], head["inflight"], bool(args.graph)),
        "roofline": roofline_of(head, peaks, traffic),
        "e2e": head.get("e2e"),
        **({"e2e_resident": head["e2e_resident"]} if "e2e_resident" in head else {}),
        "gpu_launches": head["steps"] * head["launches_per_step"],
        "serial": {"ms_per_step": head["serial_ms_per_step"], "value": world * head["batch"] / (head["serial_ms_per_step"] * 1e-3),
                   "note": "the same K steps issued on ONE stream (a step starts when the previous one has finished)"},
        **({"parts": head["parts"]} if "parts" in head else {}),
        **({"verify": head["verify"]} if "verify" in head else {}),
        "clocks": head["clocks"],
    }
    if per_config:
        pc = {}
        for k, r in per_config.items():
            if k == "datamodule":
                line["datamodule"] = r
                continue
            if "ms_per_step" not in r or k == "5":
                pc[k] = r
                continue
            pc[k] = {"workload": r["name"], "batch_per_gpu": r["batch"], "inflight": r["inflight"], "value": r["value"], "ms_per_step": r["ms_per_step"],
                     "steps": r["steps"], "serial_ms_per_step": r["serial_ms_per_step"],
                     "roofline": {k2: v for k2, v in roofline_of(r, peaks, traffic).items() if k2 in ("achieved", "frac", "traffic", "algorithmic_bytes_per_image")},
                     "e2e": r.get("e2e"), "e2e_resident": r.get("e2e_resident"), "verify": r.get("verify"), "gpu_launches": r["steps"] * r["launches_per_step"],
                     "clocks": r["clocks"]}
        line["per_config"] = pc
    if world == 1 and not args.no_cpu:
        cores = host_cores()
        big = args.config == 4               # ~1.5 s per image and core (obj_mask of a 1024x1024 image on the CPU)
        v, dt, sample = cpu_oracle_images_per_s(args.config, head["batch"], (1 if big else 4) * cores, cores, target_s=12.0)
        v1, dt1, s1 = cpu_oracle_images_per_s(args.config, head["batch"], 2 if big else 16, 1, target_s=4.0)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{sample} images cycling through the same batch, {cores} processes x 1 torch/BLAS thread "
                                          f"({dt:.1f} s); 1 process: {v1:.1f} images/s over {s1} images",
                                "value_1core": v1}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    bad = sum(int((r.get("verify") or {}).get("mismatch", 0)) for r in [head] + [r for r in per_config.values() if isinstance(r, dict)])
    if bad:
        sys.stderr.write(f"bench.py: {bad} image(s) of the timed buffers differ from the oracle\n")
        sys.exit(3)


if __name__ == "__main__":
    main()
