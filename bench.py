#!/usr/bin/env python
"""bench.py - throughput of the tensor-marshalling hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--also LIST]

One JSON line on stdout (rank 0). A "step" is one pass of the hot path over one batch of
synthetic input. The primary line is cfg3 (BASELINE.json configs[2], the configuration the
images/s metric is quoted on): 256 synthetic 1920x1080 RGB uint8 images per GPU -> 224x224 fp32
NCHW, VGG mean subtraction, written into a CUDA shared-memory region (weak scaling).

Every other BASELINE config is measured in the same run and reported under "also", at every N,
each with its own `roofline` (and `cpu_baseline` on rank 0 at N=1):
  cfg1_latency     one 640x480 image through the preprocess() drop-in (median us per call)
  cfg2_serialize   1M variable-length strings -> region            (sharded by element range)
  cfg4_deserialize 16M strings (~2.1 GB) region -> offsets+payload  (sharded by region)
  cfg5_mixed       8192 mixed-resolution images -> fp16 NHWC        (sharded by image, strong)
  bf16             64M float32 <-> bfloat16 wire bytes
  region_copy      set/get of a region from pageable numpy memory, staged vs plain
  long_elements    deserialize with 1 KiB / 64 KiB / 1 MiB elements (the repair path)
`--also a,b,c` selects a subset, `--also none` skips them all.

Multi-GPU: launched under torchrun (one rank per GPU); work shards by image / element range /
region, every rank fills its own region, there is no data-path collective. NCCL is used only
for the barrier and the max-over-ranks of the device-timed regions.
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
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CFG3 = {"n": 256, "W": 1920, "H": 1080, "c": 3, "h": 224, "w": 224, "format": 2, "dtype": "FP32",
        "scaling": "VGG"}
CFG3_ALGO_BYTES_PER_IMAGE = 1920 * 1080 * 3 + 3 * 224 * 224 * 4  # 6 822 912
CFG5_MENU = [(640, 480), (800, 600), (1024, 768), (1280, 720), (1920, 1080), (500, 375), (375, 500),
             (224, 224), (150, 100)]  # SURVEY.md 8d (W, H)
ALSO_ALL = ["cfg1_latency", "cfg2_serialize", "cfg4_deserialize", "cfg5_mixed", "bf16", "region_copy",
            "long_elements"]


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full
    capture (profiles/traffic.json), or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def roofline(algo_bytes, seconds, peak, traffic_key=None, **extra):
    r = {"bound": "hbm", "achieved": algo_bytes / seconds / 1e9, "peak": peak, "unit": "GB/s",
         "frac": algo_bytes / seconds / 1e9 / peak, "traffic": ncu_traffic(traffic_key) if traffic_key else None,
         "algorithmic_bytes_per_launch": int(algo_bytes)}
    r.update(extra)
    return r


# ---- clocks sampling ------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.marks = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def mark(self):
        """Call before and after the timed region: samples between the first and the last mark
        (plus the neighbouring ones) are the ones reported as 'under load'."""
        self.marks.append(time.perf_counter())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        time.sleep(0.1)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        lo = (self.marks[0] - 0.03) if self.marks else 0.0
        hi = (self.marks[-1] + 0.03) if self.marks else float("inf")
        sm, sm_all, mx, reasons = [], [], [], set()
        for ts, ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                v, m = float(f[1]), float(f[2])
            except ValueError:
                continue
            sm_all.append(v)
            mx.append(m)
            if lo <= ts <= hi:
                sm.append(v)
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        use = sm if sm else sm_all
        return {"sm_mhz": float(np.median(use)) if use else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(use), "samples_whole_run": len(sm_all)}


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---- CPU baselines: the reference's own code where it can be imported, else its restatement ---------
def _reference_kind():
    """("reference", loader) when the unmodified reference is importable on this machine (the build
    container has /root/reference; the GPU box does not), else ("port", None): oracle/ restates it."""
    try:
        from oracle import reference_loader as rl

        if rl.available():
            return "reference", rl
    except Exception:
        pass
    return "port", None


def _cpu_pre_worker(args):
    seed, count, W, H = args
    from PIL import Image

    kind, rl = _reference_kind()
    if kind == "reference":
        pre = rl.reference_preprocess()
        call = lambda im: pre(im, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"], CFG3["scaling"], "grpc")
    else:
        from oracle import pil_bilinear as pb

        call = lambda im: pb.preprocess_pil(im, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"],
                                            CFG3["scaling"])
    rng = np.random.default_rng(seed)
    img = Image.fromarray(rng.integers(0, 256, (H, W, 3), dtype=np.uint8), "RGB")
    t0 = time.perf_counter()
    for _ in range(count):  # the same image every time (cache-warm: generous to the CPU)
        np.ascontiguousarray(call(img))
    return time.perf_counter() - t0


def cpu_preprocess_rate(images_per_worker, workers, W=1920, H=1080):
    """images/s of the reference's preprocess (PIL resize + numpy) over `workers` processes."""
    import multiprocessing as mp

    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        pool.map(_cpu_pre_worker, [(i, 1, W, H) for i in range(workers)])  # warm-up: imports, page-in
        t0 = time.perf_counter()
        pool.map(_cpu_pre_worker, [(1000 + i, images_per_worker, W, H) for i in range(workers)])
        dt = time.perf_counter() - t0
    return workers * images_per_worker / dt, dt


def cpu_images_per_worker(workers, seconds=12.0):
    """Bounded sample: about `seconds` of CPU work per step (a 1080p image costs ~11-17 ms per core)."""
    return max(4, int(seconds / 0.013 / max(workers, 1)))


def _bytes_codec():
    kind, rl = _reference_kind()
    if kind == "reference":
        u = rl.reference_utils()
        return kind, (lambda a: u.serialize_byte_tensor(a).item()), u.deserialize_bytes_tensor
    from oracle import bytes_codec as bc

    return kind, bc.serialize_elements, (lambda b: bc.deserialize_stream(b))


def _cpu_bytes_worker(args):
    seed, n = args
    _, ser, des = _bytes_codec()
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, 257, n, dtype=np.int64)
    blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8).tobytes()
    offs = np.concatenate(([0], np.cumsum(lens))).tolist()
    arr = np.empty(n, dtype=np.object_)
    arr[:] = [blob[offs[i]:offs[i + 1]] for i in range(n)]
    t0 = time.perf_counter()
    wire = ser(arr)
    t1 = time.perf_counter()
    back = des(wire)
    t2 = time.perf_counter()
    assert len(back) == n
    return t1 - t0, t2 - t1, len(wire)


def cpu_bytes_baseline(n_total=1_000_000):
    """The reference's serialize_byte_tensor / deserialize_bytes_tensor loops (utils/__init__.py:193-276)
    on n_total elements of the cfg2 recipe: one thread, then one process per core over element
    ranges. GB/s of serialized bytes."""
    import multiprocessing as mp

    kind, _, _ = _bytes_codec()
    ts, td, nb = _cpu_bytes_worker((42, n_total))
    one = {"serialize_gbs": nb / ts / 1e9, "deserialize_gbs": nb / td / 1e9, "us_per_element_ser": ts / n_total * 1e6,
           "us_per_element_des": td / n_total * 1e6}
    cores = usable_cores()
    per = max(1, n_total // cores)
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        t0 = time.perf_counter()
        res = pool.map(_cpu_bytes_worker, [(100 + i, per) for i in range(cores)])
        wall = time.perf_counter() - t0
    total = sum(r[2] for r in res)
    # every worker ran both directions back to back (plus its own input generation, not counted):
    # rate = bytes / the slowest worker's time in that direction
    all_ser = total / max(r[0] for r in res) / 1e9
    all_des = total / max(r[1] for r in res) / 1e9
    return {"kind": kind, "elements": n_total, "one_thread": one, "cores": cores,
            "all_cores": {"serialize_gbs": all_ser, "deserialize_gbs": all_des, "wall_s": wall},
            "sample": "%d elements len~U[0,256]: 1 thread on all of them, then %d processes x %d elements"
                      % (n_total, cores, per)}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on all host cores:
    imported unmodified when /root/reference exists on this machine, else its restatement in oracle/
    (PIL resize + numpy; the reference is pure Python for this path, so there is no oracle/_ref)."""
    if rank != 0:
        return
    kind, _ = _reference_kind()
    cores = usable_cores()
    workers = max(1, cores)
    per = cpu_images_per_worker(workers)
    total_steps = args.warmup + args.steps
    rates = []
    t_all = time.perf_counter()
    rate = dt = 0.0
    for s in range(total_steps):
        rate, dt = cpu_preprocess_rate(per, workers)
        if s >= args.warmup:
            rates.append((rate, dt))
        if time.perf_counter() - t_all > 200:
            break
    value = float(np.mean([r for r, _ in rates])) if rates else rate
    ms = float(np.mean([d for _, d in rates])) * 1e3 if rates else dt * 1e3
    import PIL

    sample = "%d procs x %d images of 1920x1080 per step (PIL %s resize + numpy %s), %d timed steps" % (
        workers, per, PIL.__version__, np.__version__, len(rates))
    also = {}
    if "none" not in args.also:
        try:
            also["bytes_cpu"] = cpu_bytes_baseline(200_000)
        except Exception as ex:
            also["bytes_cpu"] = {"error": repr(ex)}
    line = {
        "impl": "reference", "metric": "images/s", "value": value, "unit": "images/s", "n_gpus": args.gpus,
        "steps": len(rates), "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cfg3: 1920x1080 RGB u8 -> 224x224 fp32 NCHW VGG (bounded sample on host cores)"},
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": workers, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "also": also,
    }
    print(json.dumps(line))


# ---- GPU side --------------------------------------------------------------------------------------
class Ctx:
    """What every GPU leg needs: torch, the process group, this rank's device, peak, step counts."""

    def __init__(self, torch, dist, dev, rank, world, peak, steps, warmup):
        self.torch, self.dist, self.dev, self.rank, self.world = torch, dist, dev, rank, world
        self.peak, self.steps, self.warmup = peak, steps, warmup
        self.device = "cuda:%d" % dev
        self._flush = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def max_over_ranks(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.device)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def flush_l2(self, i=0):
        if self._flush is None:
            self._flush = self.torch.empty(256 << 20, dtype=self.torch.uint8, device=self.device)
        self._flush.fill_(i & 0xFF)

    def drop_flush(self):
        self._flush = None

    def time_calls(self, fn, steps=None, warmup=None, flush=False):
        """Median seconds per call of fn() (CUDA events on torch's current stream, which is the
        stream every library call below is launched on), max over ranks."""
        torch = self.torch
        steps = self.steps if steps is None else steps
        warmup = self.warmup if warmup is None else warmup
        times = []
        self.barrier()
        for i in range(warmup + steps):
            if flush:
                self.flush_l2(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if i >= warmup:
                times.append(e0.elapsed_time(e1) * 1e-3)
        return self.max_over_ranks(float(np.median(times)))


    def time_back_to_back(self, fns, steps, warmup):
        """Seconds per call with `steps` calls enqueued back to back between two CUDA events (how the
        primary metric is timed): launch gaps and host time between calls are not charged to the
        kernel. `fns` is cycled: callers whose working set would otherwise stay in L2 pass several
        closures over different buffers. Max over ranks."""
        torch = self.torch
        t0 = time.perf_counter()
        i = 0
        while i < warmup or time.perf_counter() - t0 < 0.05:  # short kernels: also long enough for the clocks to come up
            fns[i % len(fns)]()
            i += 1
            if i % 8 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fns[i % len(fns)]()
        e1.record()
        torch.cuda.synchronize()
        return self.max_over_ranks(e0.elapsed_time(e1) * 1e-3 / steps)


def shard(n, rank, world):
    return n * rank // world, n * (rank + 1) // world


def gbytes_recipe(n, seed):
    """SURVEY.md 8c G-BYTES recipe: lens then blob from the same generator."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, 257, n, dtype=np.int64)
    blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
    return blob, lens


def bench_cfg2(cx):
    """cfg2: G-BYTES N=1M (seed 42) device blob + int32 lengths -> a region; rank r takes the element
    range [r*N/R, (r+1)*N/R) into its own region (strong scaling)."""
    torch = cx.torch
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    N = 1_000_000
    blob, lens = gbytes_recipe(N, 42)
    e0, e1 = shard(N, cx.rank, cx.world)
    starts = np.concatenate(([0], np.cumsum(lens)))
    my_blob = blob[starts[e0]:starts[e1]]
    my_lens = lens[e0:e1].astype(np.int32)
    P, n = int(my_blob.size), int(my_lens.size)
    d_blob = torch.from_numpy(my_blob).to(cx.device)
    d_lens = torch.from_numpy(my_lens).to(cx.device)
    region = cudashm.create_shared_memory_region("bench_cfg2_rank%d" % cx.rank, P + 4 * n, cx.dev)
    # three more copies of the inputs and outputs, used in rotation: a launch then never finds its 260 MB
    # in the 126 MB L2 although the launches run back to back
    rot = [(d_blob.clone(), d_lens.clone(), torch.empty(P + 4 * n, dtype=torch.uint8, device=cx.device)) for _ in range(3)]
    try:
        t_iso = cx.time_calls(lambda: u.serialize_to_device(d_blob, d_lens, region, 0, cx.dev, sync=False),
                              steps=min(cx.steps, 30), warmup=min(cx.warmup, 5), flush=True)
        fns = [lambda: u.serialize_to_device(d_blob, d_lens, region, 0, cx.dev, sync=False)]
        fns += [(lambda b, l, o: (lambda: u.serialize_to_device(b, l, o, 0, cx.dev, sync=False)))(b, l, o) for b, l, o in rot]
        t = cx.time_back_to_back(fns, steps=4 * max(1, min(cx.steps, 32) // 4), warmup=4)
        ok = None
        if cx.rank == 0:  # parity of rank 0's shard against the oracle (checker only)
            from oracle import clib

            want = clib.bytes_serialize(my_blob, my_lens.astype(np.int64))
            got = cudashm.get_contents_as_numpy(region, np.uint8, [P + 4 * n])
            ok = bool(np.array_equal(got, want))
        tot_P, tot_n = int(blob.size), N
        res = {"metric": "BYTES serialize GB/s", "value": (tot_P + 4 * tot_n) / t / 1e9,
               "unit": "GB/s of serialized output", "us_per_launch": t * 1e6, "elements": N, "n_gpus": cx.world,
               "scaling": "strong", "bit_exact_vs_oracle_rank0": ok,
               "roofline": roofline(2 * (P + 4 * n), t, cx.peak, "k_bytes_serialize_cfg2",
                                    note="per rank: this rank's algorithmic bytes / max-over-ranks time"),
               "config": {"workload": "cfg2: 1M strings len~U[0,256] (G-BYTES seed 42), device blob+int32 lengths -> "
                                      "cudashm region, element range per rank",
                          "l2": "four input/output sets (1 GB) used in rotation, launches back to back between two events",
                          "includes": "workspace clear + scan+scatter kernel (one launch)"},
               "us_per_call_isolated": t_iso * 1e6,
               "isolated": "one call at a time, events around each, 256 MiB L2 flush write in between (launch gap and host time included)"}
        # the same kernel at cfg4's size, so that latency/tail effects of a 40 us problem are separated
        if cx.world == 1:
            g = torch.Generator(device=cx.device).manual_seed(7)
            N2 = 16_000_000
            l2 = torch.randint(0, 257, (N2,), dtype=torch.int32, device=cx.device, generator=g)
            P2 = int(l2.sum(dtype=torch.int64).item())
            b2 = torch.randint(0, 256, (P2,), dtype=torch.uint8, device=cx.device, generator=g)
            o2 = torch.empty(P2 + 4 * N2, dtype=torch.uint8, device=cx.device)
            t2 = cx.time_back_to_back([lambda: u.serialize_to_device(b2, l2, o2, 0, cx.dev, sync=False)], steps=5, warmup=2)
            res["at_16M_elements"] = {"ms_per_launch": t2 * 1e3, "roofline": roofline(2 * (P2 + 4 * N2), t2, cx.peak)}
            del l2, b2, o2
        return res
    finally:
        del rot
        cudashm.destroy_shared_memory_region(region)


def bench_cfg4(cx):
    """cfg4: 16M strings (seed 43, ~2.1 GB serialized) in regions -> int64 offsets + compact payload.
    Sharded by region: rank r holds the elements [r*N/R, (r+1)*N/R) in its own region."""
    torch = cx.torch
    from triton_client_b200 import _lib
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    N = 16_000_000
    rng = np.random.default_rng(43)
    lens = rng.integers(0, 257, N, dtype=np.int64)
    e0, e1 = shard(N, cx.rank, cx.world)
    my_lens = lens[e0:e1]
    P, n = int(my_lens.sum()), int(my_lens.size)
    # payload bytes: generated on the device (2 GB from numpy's generator would cost seconds per rank)
    g = torch.Generator(device=cx.device).manual_seed(43 + cx.rank)
    d_blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device=cx.device, generator=g)
    d_lens = torch.from_numpy(my_lens.astype(np.int32)).to(cx.device)
    region = cudashm.create_shared_memory_region("bench_cfg4_rank%d" % cx.rank, P + 4 * n, cx.dev)
    try:
        _, tot = u.serialize_to_device(d_blob, d_lens, region, 0, cx.dev)
        assert tot == P + 4 * n
        prefix_ok = None
        if cx.rank == 0:  # the stream the decoder reads is the oracle's, checked on a 200k-element prefix
            from oracle import clib

            m = min(n, 200_000)
            pb = int(my_lens[:m].sum())
            want = clib.bytes_serialize(d_blob[:pb].cpu().numpy(), my_lens[:m])
            got = cudashm.get_contents_as_numpy(region, np.uint8, [pb + 4 * m])
            prefix_ok = bool(np.array_equal(got, want))
        out = {}

        def call():
            out["r"] = u.deserialize_from_device(region, tot, -1, 0, cx.dev)

        launches0 = _lib.kernel_launch_count()
        t = cx.time_calls(call, steps=max(3, min(cx.steps, 10)), warmup=2)
        launches = (_lib.kernel_launch_count() - launches0) // (max(3, min(cx.steps, 10)) + 2)
        offsets, payload = out["r"]
        ok = bool(offsets.numel() == n + 1 and payload.numel() == P and torch.equal(payload, d_blob)
                  and torch.equal(offsets[1:] - offsets[:-1], d_lens.to(torch.int64)))
        tot_all = cx.sum_over_ranks(tot)
        algo = tot + P + 8 * (n + 1)
        return {"metric": "BYTES deserialize GB/s", "value": tot_all / t / 1e9, "unit": "GB/s of serialized input",
                "ms_per_call": t * 1e3, "elements": N, "n_gpus": cx.world, "scaling": "strong",
                "decoded_equals_input": ok, "stream_prefix_equals_oracle_rank0": prefix_ok, "launches_per_call": launches,
                "roofline": roofline(algo, t, cx.peak, "deserialize_cfg4_all_kernels",
                                     note="per rank: this rank's algorithmic bytes / max-over-ranks time; all kernels of one call"),
                "config": {"workload": "cfg4: 16M strings len~U[0,256] (lengths: numpy seed 43; ~2.1 GB stream) in cudashm "
                                       "regions -> int64 offsets + compact payload on the device, one region per rank",
                           "l2": "input >> L2 at N=1", "includes": "output/workspace allocation, all kernels, the small "
                           "D2H header reads the call makes"}}
    finally:
        cudashm.destroy_shared_memory_region(region)


def cfg5_sizes(n=8192):
    idx = np.random.default_rng(1234).choice(len(CFG5_MENU), size=n)
    return [CFG5_MENU[i] for i in idx]


def bench_cfg5(cx):
    """cfg5: 8192 mixed-resolution images (SURVEY.md 8d menu, default_rng(1234).choice) -> 224x224
    fp16 NHWC INCEPTION; rank r takes images [r*8192/R, (r+1)*8192/R) into its own region."""
    torch = cx.torch
    from triton_client_b200.image_client import PreprocessPlan
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    sizes = cfg5_sizes()
    i0, i1 = shard(len(sizes), cx.rank, cx.world)
    mine = sizes[i0:i1]
    total_src = sum(w * h * 3 for w, h in mine)
    al = lambda x: (x + 255) & ~255  # every image starts on a 256-byte boundary, as separate allocations would
    offs, off = [], 0
    for w, h in mine:
        offs.append(off)
        off = al(off + w * h * 3)
    g = torch.Generator(device=cx.device).manual_seed(5000 + cx.rank)
    pixels = torch.randint(0, 256, (off,), dtype=torch.uint8, device=cx.device, generator=g)
    base = pixels.data_ptr()
    ptrs = np.array([base + o for o in offs], dtype=np.uint64)  # unchanged between steps: the plan re-uploads nothing
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    plan = PreprocessPlan([(w, h, 3) for w, h in mine], 1, "FP16", 3, 224, 224, "INCEPTION", device=cx.dev)
    out_bytes = len(mine) * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("bench_cfg5_rank%d" % cx.rank, out_bytes, cx.dev)
    try:
        plan.run(ptrs, region._base_addr, out_bytes)  # the plan is created on its first run
        torch.cuda.synchronize()
        plan_ms = cx.max_over_ranks((time.perf_counter() - t0) * 1e3)
        info = plan.info()
        t = cx.time_calls(lambda: plan.run(ptrs, region._base_addr, out_bytes), steps=min(cx.steps, 20),
                          warmup=min(cx.warmup, 3))
        ok = None
        if cx.rank == 0:  # one image of every size on rank 0 against the oracle
            from oracle import clib
            from oracle import pil_bilinear as pb

            lut = pb.scaling_lut("FP16", 3, "INCEPTION")
            got = cudashm.get_contents_as_numpy(region, np.float16, [len(mine), 224, 224, 3])
            seen, ok = set(), True
            for k, (w, h) in enumerate(mine):
                if (w, h) not in seen:
                    seen.add((w, h))
                    src = pixels[offs[k]:offs[k] + w * h * 3].cpu().numpy().reshape(h, w, 3)
                    u8 = clib.resize_bilinear_u8(src, 224, 224)
                    want = np.stack([lut[ch][u8[:, :, ch]] for ch in range(3)], axis=-1)
                    ok = ok and bool(np.array_equal(got[k].view(np.uint16), want.view(np.uint16)))
        algo_mine = total_src + len(mine) * 224 * 224 * 3 * 2
        n_all = len(sizes)
        return {"metric": "images/s", "value": n_all / t, "unit": "images/s", "ms_per_batch": t * 1e3, "images": n_all,
                "n_gpus": cx.world, "scaling": "strong", "plan_build_ms": plan_ms,
                "fused_images_this_rank": info["fused_images"], "images_this_rank": len(mine),
                "parity_one_image_per_size_rank0": ok,
                "roofline": roofline(algo_mine, t, cx.peak, None,
                                     note="per rank: sum(W*H*3) + n*301056 bytes of this rank / max-over-ranks time"),
                "config": {"workload": "cfg5: 8192 images, sizes drawn by default_rng(1234).choice from the 9-size menu, "
                                       "uniform pixels -> (n,224,224,3) fp16 NHWC INCEPTION, one region per rank",
                           "l2": "input %.1f GB per rank >> L2" % (total_src / 1e9),
                           "plan_build_ms": "PreprocessPlan construction + first run (coefficient tables, descriptors, "
                                            "work partition, uploads), not part of the timed steps"}}
    finally:
        plan.close()
        cudashm.destroy_shared_memory_region(region)


def bench_cfg1(cx, cpu):
    """cfg1: one synthetic 640x480 RGB image (seed 0) -> (3,224,224) f32 INCEPTION through the
    preprocess() drop-in, host PIL image in, numpy array out. Launch/latency-bound: microseconds."""
    if cx.rank != 0:
        return None
    from PIL import Image

    from triton_client_b200.image_client import preprocess

    a = np.random.default_rng(0).integers(0, 256, (480, 640, 3), dtype=np.uint8)
    img = Image.fromarray(a, "RGB")
    for _ in range(10):
        out = preprocess(img, 2, "FP32", 3, 224, 224, "INCEPTION", "grpc")
    ts = []
    for _ in range(200):
        t0 = time.perf_counter()
        out = preprocess(img, 2, "FP32", 3, 224, 224, "INCEPTION", "grpc")
        ts.append(time.perf_counter() - t0)
    import hashlib

    sha = hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest()[:16]
    res = {"metric": "preprocess() latency", "value": float(np.median(ts)) * 1e6, "unit": "us per call (median of 200)",
           "higher_is_better": False, "p10_us": float(np.percentile(ts, 10)) * 1e6,
           "p90_us": float(np.percentile(ts, 90)) * 1e6,
           "golden_sha256_prefix_matches": sha == "b4ab8af0fee98c6d",  # SURVEY.md 8c seed0 golden
           "config": {"workload": "cfg1: PIL RGB 640x480 (seed 0) -> (3,224,224) f32 INCEPTION, pageable H2D + fused kernel "
                                  "+ D2H per call; launch/latency-bound, no roofline (1.5 MB of traffic = 0.23 us of HBM time)"}}
    if cpu:
        kind, rl = _reference_kind()
        if kind == "reference":
            pre = rl.reference_preprocess()
            call = lambda: np.ascontiguousarray(pre(img, 2, "FP32", 3, 224, 224, "INCEPTION", "grpc"))
        else:
            from oracle import pil_bilinear as pb

            call = lambda: np.ascontiguousarray(pb.preprocess_pil(img, 2, "FP32", 3, 224, 224, "INCEPTION"))
        for _ in range(5):
            call()
        tc = []
        for _ in range(100):
            t0 = time.perf_counter()
            call()
            tc.append(time.perf_counter() - t0)
        res["cpu_baseline"] = {"value": float(np.median(tc)) * 1e6, "unit": "us per call (median of 100)", "cores": 1,
                               "kind": kind, "sample": "the same call through the reference's preprocess (PIL + numpy)"}
    return res


def bench_bf16(cx):
    """f3: 64M float32 -> bfloat16 wire bytes into a region, and back (6 bytes of traffic per element)."""
    torch = cx.torch
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    n = 64 << 20
    vals = torch.randn(n, device=cx.device)
    region = cudashm.create_shared_memory_region("bench_bf16_rank%d" % cx.rank, 2 * n, cx.dev)
    back = torch.empty(n, dtype=torch.float32, device=cx.device)
    try:
        ts = cx.time_back_to_back([lambda: u.serialize_bf16_to_device(vals, region, 0, cx.dev, sync=False)],
                                  steps=min(cx.steps, 20), warmup=3)
        td = cx.time_back_to_back([lambda: u.deserialize_bf16_from_device(region, 2 * n, 0, back, cx.dev, sync=False)],
                                  steps=min(cx.steps, 20), warmup=3)
        ok = None
        if cx.rank == 0:
            from oracle import bytes_codec as bc

            m = 1 << 20
            wire = cudashm.get_contents_as_numpy(region, np.uint8, [2 * m])
            ok = bool(wire.tobytes() == bc.bf16_serialize(vals[:m].cpu().numpy())
                      and back[:m].cpu().numpy().tobytes() == bc.bf16_deserialize(wire.tobytes()).tobytes())
        return {"metric": "BF16 codec GB/s", "value": cx.world * 6 * n / ts / 1e9, "unit": "GB/s of traffic (serialize)",
                "elements_per_gpu": n, "n_gpus": cx.world, "scaling": "weak", "serialize_us": ts * 1e6,
                "deserialize_us": td * 1e6, "bit_exact_vs_oracle_rank0": ok,
                "roofline": roofline(6 * n, ts, cx.peak, None, kernel="k_bf16_pack"),
                "roofline_deserialize": roofline(6 * n, td, cx.peak, None, kernel="k_bf16_expand"),
                "config": {"workload": "64 Mi float32 (256 MiB) -> 128 MiB of bfloat16 wire bytes in a cudashm region, and back",
                           "l2": "384 MiB of traffic per launch >> L2; launches enqueued back to back between two events"}}
    finally:
        cudashm.destroy_shared_memory_region(region)


def bench_region_copy(cx):
    """a7/a8 at the PCIe edge: set_shared_memory_region / get_contents_as_numpy from ordinary numpy
    memory, through the library's pinned staging ring and (TCB_STAGING=0 in a child process) the
    reference-style plain pageable copy."""
    if cx.rank != 0:
        return None
    code = r'''
import json, sys, time
import numpy as np
sys.path.insert(0, %r)
from triton_client_b200.utils import cuda_shared_memory as cudashm
res = {}
for name, nbytes in (("154MB", 154140672), ("2.1GB", 2112000000)):
    a = np.random.default_rng(1).integers(0, 256, nbytes, dtype=np.uint8)
    r = cudashm.create_shared_memory_region("copy", nbytes, 0)
    cudashm.set_shared_memory_region(r, [a])
    ts, tg = [], []
    for _ in range(3):
        t0 = time.perf_counter(); cudashm.set_shared_memory_region(r, [a]); ts.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); b = cudashm.get_contents_as_numpy(r, np.uint8, [nbytes]); tg.append(time.perf_counter() - t0)
    res[name] = {"set_gbs": nbytes / min(ts) / 1e9, "get_gbs": nbytes / min(tg) / 1e9, "get_first_call_gbs": nbytes / tg[0] / 1e9,
                 "roundtrip_equal": bool(np.array_equal(a, b))}
    cudashm.destroy_shared_memory_region(r)
    del a, b
print(json.dumps(res))
''' % ROOT
    out = {}
    for label, env in (("staged", {}), ("staged_no_result_pool", {"TCB_HOST_POOL_BYTES": "0"}),
                       ("plain_pageable", {"TCB_STAGING": "0", "TCB_HOST_POOL_BYTES": "0"})):
        e = dict(os.environ)
        e.update(env)
        e["CUDA_VISIBLE_DEVICES"] = os.environ.get("CUDA_VISIBLE_DEVICES", "") or str(cx.dev)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
            e.pop(k, None)
        try:
            p = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=240)
            out[label] = json.loads(p.stdout.strip().splitlines()[-1])
        except Exception as ex:
            out[label] = {"error": repr(ex)[:300]}
    try:
        out["speedup_set_2.1GB"] = out["staged"]["2.1GB"]["set_gbs"] / out["plain_pageable"]["2.1GB"]["set_gbs"]
        out["speedup_get_2.1GB"] = out["staged"]["2.1GB"]["get_gbs"] / out["plain_pageable"]["2.1GB"]["get_gbs"]
    except Exception:
        pass
    out["config"] = {"workload": "set_shared_memory_region / get_contents_as_numpy of 154 MB and 2.1 GB uint8 arrays in "
                                 "pageable numpy memory (best of 3 wall-clock calls, each ends synchronised)",
                     "staged": "set: 4 x 4 MiB pinned ring + 7 host copy threads (staging.cu); get: the result array lives in a "
                               "recycled page-locked buffer (first call allocates it: get_first_call_gbs)",
                     "staged_no_result_pool": "get into a fresh pageable numpy array through the ring (first-touch page faults dominate)",
                     "plain_pageable": "cudaMemcpyAsync on the numpy buffer, as the reference does"}
    return out


def bench_long_elements(cx):
    """Deserialize with elements longer than the 512-byte entry window: these streams take
    the repair path. 256 MiB of stream each."""
    if cx.rank != 0:
        return None
    torch = cx.torch
    from triton_client_b200 import utils as u

    res = {}
    for name, elen in (("1KiB", 1024), ("64KiB", 65536), ("1MiB", 1 << 20)):
        n = (256 << 20) // elen
        lens = torch.full((n,), elen, dtype=torch.int32, device=cx.device)
        blob = torch.randint(0, 256, (n * elen,), dtype=torch.uint8, device=cx.device)
        stream, tot = u.serialize_to_device(blob, lens)
        out = {}

        def call():
            out["r"] = u.deserialize_from_device(stream, tot)

        t = cx.time_calls(call, steps=3, warmup=1)
        offsets, payload = out["r"]
        ok = bool(offsets.numel() == n + 1 and torch.equal(payload, blob))
        algo = tot + n * elen + 8 * (n + 1)
        res[name] = {"elements": n, "ms_per_call": t * 1e3, "gbs_of_stream": tot / t / 1e9, "decoded_equals_input": ok,
                     "roofline": roofline(algo, t, cx.peak)}
        del lens, blob, stream, offsets, payload, out
    res["config"] = {"workload": "256 MiB streams of equal-length elements (1 KiB, 64 KiB, 1 MiB), random payload"}
    return res


def bench_bytes_api(cx):
    """API-level BYTES numbers (np.object_ array in -> region, region -> np.object_ array out): the
    host-side object work the kernels cannot remove, reported separately."""
    if cx.rank != 0:
        return None
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    N = 1_000_000
    blob, lens = gbytes_recipe(N, 42)
    offs = np.concatenate(([0], np.cumsum(lens))).tolist()
    b = blob.tobytes()
    arr = np.empty(N, dtype=np.object_)
    arr[:] = [b[offs[i]:offs[i + 1]] for i in range(N)]
    total = len(b) + 4 * N
    region = cudashm.create_shared_memory_region("bench_api", total, cx.dev)
    try:
        u.serialize_byte_tensor_to(region, arr)
        ts, td = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            nb = u.serialize_byte_tensor_to(region, arr)
            ts.append(time.perf_counter() - t0)
            t0 = time.perf_counter()
            back = cudashm.get_contents_as_numpy(region, np.object_, [N])
            td.append(time.perf_counter() - t0)
        ok = bool(nb == total and back.shape == (N,) and back[0] == arr[0] and back[N - 1] == arr[N - 1]
                  and all(back[i] == arr[i] for i in range(0, N, 997)))
        # a request's handful of strings: latency of the host-to-host drop-ins, next to the reference's loops
        from oracle import bytes_codec as obc

        def med_us(fn, reps):
            for _ in range(10):
                fn()
            xs = []
            for _ in range(reps):
                t0 = time.perf_counter()
                fn()
                xs.append(time.perf_counter() - t0)
            return float(np.median(xs) * 1e6)

        small = {}
        for k in (1, 16, 256):
            a = np.empty(k, dtype=np.object_)
            for i in range(k):
                a[i] = b"0.%06d:%d:label number %d" % (i * 7919 % 1000000, i, i)
            ser = u.serialize_byte_tensor(a).item()
            small[str(k)] = {"serialize_us": med_us(lambda: u.serialize_byte_tensor(a), 200),
                             "deserialize_us": med_us(lambda: u.deserialize_bytes_tensor(ser), 200),
                             "cpu_serialize_us": med_us(lambda: obc.serialize_elements(a), 50),
                             "cpu_deserialize_us": med_us(lambda: obc.deserialize_stream(ser), 50),
                             "equal": bool(ser == obc.serialize_elements(a)
                                           and u.deserialize_bytes_tensor(ser).tolist() == obc.deserialize_stream(ser))}
        return {"object_array_to_region_ms": min(ts) * 1e3, "region_to_object_array_ms": min(td) * 1e3,
                "serialize_gbs": total / min(ts) / 1e9, "deserialize_gbs": total / min(td) / 1e9, "elements": N,
                "sampled_equal": ok,
                "small_tensors": small,
                "small_tensors_note": "median us per call of serialize_byte_tensor / deserialize_bytes_tensor (host bytes in, "
                                      "host bytes out) for 1 / 16 / 256 short strings; cpu_* = the reference's Python loops "
                                      "(oracle port). Launch-latency-bound: one kernel launch and one synchronisation each",
                "config": {"workload": "cfg2 recipe as a 1-D np.object_ array of 1M bytes objects -> serialize_byte_tensor_to("
                                       "region); get_contents_as_numpy(region, np.object_, [1M]); wall clock, best of 3"}}
    finally:
        cudashm.destroy_shared_memory_region(region)


def bind_near_gpu(dev):
    """Pin this process to the cores `nvidia-smi topo -m` lists for its GPU (before any pinned
    allocation). Returns a short description for the JSON line."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        for ln in out.splitlines():
            f = ln.split()
            if f and f[0] == "GPU%d" % dev:
                # ... CPU Affinity column looks like "0-31" or "0-15,64-79"
                for tok in f[1:]:
                    if tok and tok[0].isdigit() and all(c in "0123456789-," for c in tok):
                        cores = set()
                        for part in tok.split(","):
                            a, _, b = part.partition("-")
                            cores.update(range(int(a), int(b or a) + 1))
                        allowed = cores & set(os.sched_getaffinity(0))
                        if allowed:
                            os.sched_setaffinity(0, allowed)
                            return "cores %s (nvidia-smi topo -m)" % tok
        return "no affinity column for GPU%d; left unchanged" % dev
    except Exception as ex:
        return "unchanged (%s)" % type(ex).__name__


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    # 100 steps = 43 ms of back-to-back launches: long enough for three clock samples, short enough that the
    # board does not reach its power cap (measured: 0.4275 ms/step at 50-100 steps, 0.437 at 500 and
    # 0.447 at 1000 with sw_power_cap active)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--also", default="all", help="comma list of side benchmarks (%s), 'all' or 'none'" % ",".join(ALSO_ALL))
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    ap.add_argument("--no-e2e", action="store_true", help="kernel-only run (experiments; the e2e object is then null)")
    args = ap.parse_args()
    args.also = ALSO_ALL if args.also == "all" else [x for x in args.also.split(",") if x]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3
    t_start = time.perf_counter()

    # CPU baselines first (rank 0, N=1), before CUDA is initialised in this process: the worker
    # pools are forked and only ever touch PIL / numpy.
    cpu = cpu_bytes = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = usable_cores()
        per = cpu_images_per_worker(cores)
        rate, dt = cpu_preprocess_rate(per, cores)
        import PIL

        kind, _ = _reference_kind()
        cpu = {"value": rate, "unit": "images/s", "cores": cores, "kind": kind,
               "sample": "%d procs x %d images of 1920x1080 (the reference's preprocess: Pillow %s resize + numpy %s), "
                         "%.1f s" % (cores, per, PIL.__version__, np.__version__, dt)}
        if "none" not in args.also and any(a.startswith("cfg2") or a.startswith("cfg4") for a in args.also):
            try:
                cpu_bytes = cpu_bytes_baseline(1_000_000)
            except Exception as ex:
                cpu_bytes = {"error": repr(ex)}

    dev = local_rank
    affinity = bind_near_gpu(dev)

    import torch
    import torch.distributed as dist

    from triton_client_b200 import _lib
    from triton_client_b200.image_client import HostBatchPipeline, PreprocessPlan
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % dev))
    peak, peak_src = measured_peak_hbm()
    cx = Ctx(torch, dist, dev, rank, world, peak, args.steps, args.warmup)
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()  # before warm-up, so that even a 10 ms timed region has samples around it

    n, H, W = CFG3["n"], CFG3["H"], CFG3["W"]
    g = torch.Generator(device=cx.device).manual_seed(1000 + rank)
    batch = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=cx.device, generator=g)
    plan = PreprocessPlan([(W, H, 3)] * n, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"],
                          CFG3["scaling"], device=dev)
    out_bytes = n * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("bench_out_rank%d" % rank, out_bytes, dev)
    ptrs = [batch.data_ptr() + i * H * W * 3 for i in range(n)]

    def step():
        plan.run(ptrs, region._base_addr, out_bytes)

    # ---- device-resident timing (value, roofline) -----------------------------------------------
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    cx.barrier()
    launches0 = _lib.kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    sampler.mark()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    sampler.mark()
    # stop polling here: nvidia-smi every 20 ms takes driver locks that the copy-heavy legs below
    # (several hundred cudaMemcpyAsync / event calls per pass) would be timed against
    clocks = sampler.stop() if rank == 0 else None
    cx.barrier()
    launches = _lib.kernel_launch_count() - launches0
    t_local = e0.elapsed_time(e1) * 1e-3
    t_total = cx.max_over_ranks(t_local)
    ms_per_step = t_total / args.steps * 1e3
    value = world * n * args.steps / t_total
    info = plan.info()

    # parity spot check inside the bench (oracle as the checker only): three images on every rank
    from oracle import clib
    from oracle import pil_bilinear as pb

    got = cudashm.get_contents_as_numpy(region, np.float32, [n, 3, 224, 224])
    lut = pb.scaling_lut("FP32", 3, "VGG")
    parity_local = True
    for k in (0, 5 + rank, n - 1):
        u8 = clib.resize_bilinear_u8(batch[k].cpu().numpy(), 224, 224)
        parity_local = parity_local and bool(np.array_equal(got[k], np.stack([lut[ch][u8[:, :, ch]] for ch in range(3)])))
    parity = cx.sum_over_ranks(1.0 if parity_local else 0.0) == world
    del got

    e2e_value = e2e_ok = h2d_ceiling_gbs = bidir_ceiling_gbs = pageable = None
    e2e_steps, in_bytes = 0, n * H * W * 3
    if args.no_e2e:
        del batch
        torch.cuda.empty_cache()
    else:
        # ---- end to end: host buffers in, result read back, through the public API -------------------
        h_batch = torch.empty((n, H, W, 3), dtype=torch.uint8, pin_memory=True)
        h_batch.copy_(batch)
        h_out = torch.empty(out_bytes, dtype=torch.uint8, pin_memory=True)
        in_bytes = n * H * W * 3

        # (a) what the box gives: a bare pinned -> device copy of the same bytes on all ranks at once
        stage = torch.empty(in_bytes, dtype=torch.uint8, device=cx.device)
        cur = torch.cuda.current_stream(dev).cuda_stream

        def bare_copy():
            _lib.check(_lib.lib.TcbMemcpy(dev, stage.data_ptr(), h_batch.data_ptr(), in_bytes, cur, 0))

        t_copy = cx.time_calls(bare_copy, steps=5, warmup=2)
        h2d_ceiling_gbs = world * in_bytes / t_copy / 1e9
        # ... and with the result's read-back running against it on a second stream, which is what the
        # pipeline does: both PCIe directions at once, on all ranks at once
        s_a, s_b = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        d_res = torch.empty(out_bytes, dtype=torch.uint8, device=cx.device)

        def bare_both():
            here = torch.cuda.current_stream(dev)
            s_a.wait_stream(here)
            s_b.wait_stream(here)
            _lib.check(_lib.lib.TcbMemcpy(dev, stage.data_ptr(), h_batch.data_ptr(), in_bytes, s_a.cuda_stream, 0))
            _lib.check(_lib.lib.TcbMemcpy(dev, h_out.data_ptr(), d_res.data_ptr(), out_bytes, s_b.cuda_stream, 0))
            here.wait_stream(s_a)
            here.wait_stream(s_b)

        t_both = cx.time_calls(bare_both, steps=5, warmup=2)
        bidir_ceiling_gbs = world * in_bytes / t_both / 1e9
        del stage, d_res

        pipe = HostBatchPipeline(n, W, H, 3, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"],
                                 CFG3["scaling"], device=dev, chunks=int(os.environ.get("TCB_BENCH_CHUNKS", "8")))

        def e2e_step():
            pipe.run(h_batch.data_ptr(), region._base_addr, out_bytes, h_out.data_ptr())

        e2e_steps = max(3, min(args.steps, 10))
        for _ in range(3):
            e2e_step()
        torch.cuda.synchronize()
        cx.barrier()
        e0.record()
        for _ in range(e2e_steps):
            e2e_step()
        e1.record()
        torch.cuda.synchronize()
        cx.barrier()
        t_e2e = cx.max_over_ranks(e0.elapsed_time(e1) * 1e-3)
        e2e_value = world * n * e2e_steps / t_e2e
        # the bytes that came back over PCIe are the ones the device-resident run produced
        ref_out = torch.empty(out_bytes, dtype=torch.uint8, pin_memory=True)
        step()
        _lib.check(_lib.lib.TcbRegionGet(region._c_handle, 0, ref_out.data_ptr(), out_bytes,
                                         torch.cuda.current_stream(dev).cuda_stream))
        e2e_ok = cx.sum_over_ranks(1.0 if torch.equal(ref_out, h_out) else 0.0) == world

        # (c) the same pipeline from ordinary pageable memory (what the reference API hands over)
        pageable = None
        try:
            src_np = np.empty((n, H, W, 3), dtype=np.uint8)
            np.copyto(src_np, h_batch.numpy())
            out_np = np.empty(out_bytes, dtype=np.uint8)
            pipe.run_pageable(src_np, region._base_addr, out_bytes, out_np)
            cx.barrier()
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                pipe.run_pageable(src_np, region._base_addr, out_bytes, out_np)
            torch.cuda.synchronize()
            t_pg = cx.max_over_ranks(time.perf_counter() - t0)
            pg_ok = cx.sum_over_ranks(1.0 if np.array_equal(out_np, ref_out.numpy()) else 0.0) == world
            pageable = {"value": world * n * reps / t_pg, "unit": "images/s", "matches_device_run": pg_ok,
                        "path": "HostBatchPipeline.run_pageable: numpy array -> the library's pinned staging ring (parallel host copy) -> "
                                "H2D | kernel | D2H -> numpy array (wall clock, %d passes)" % reps}
            del src_np, out_np
        except Exception as ex:
            pageable = {"error": repr(ex)[:300]}
        pipe.close()
        del h_batch, pipe, ref_out, h_out, batch
        torch.cuda.empty_cache()

    # ---- the other BASELINE configs ---------------------------------------------------------------
    also = {}
    legs = {"cfg2_serialize": lambda: bench_cfg2(cx), "cfg4_deserialize": lambda: bench_cfg4(cx),
            "cfg5_mixed": lambda: bench_cfg5(cx), "cfg1_latency": lambda: bench_cfg1(cx, not args.no_cpu and world == 1),
            "bf16": lambda: bench_bf16(cx),
            "region_copy": (lambda: bench_region_copy(cx)) if world == 1 else (lambda: None),
            "long_elements": (lambda: bench_long_elements(cx)) if world == 1 else (lambda: None)}
    for name in args.also:
        if name not in legs:
            continue
        try:
            r = legs[name]()
            if r is not None and rank == 0:
                also[name] = r
        except Exception as ex:  # the primary line must still be printed
            if world > 1:
                raise  # a rank that skips a leg would leave the others in its collectives
            also[name] = {"error": repr(ex)[:400]}
        cx.drop_flush()
        torch.cuda.empty_cache()
    if rank == 0 and world == 1 and "cfg2_serialize" in args.also:
        try:
            also["bytes_api_level"] = bench_bytes_api(cx)
        except Exception as ex:
            also["bytes_api_level"] = {"error": repr(ex)[:400]}
    if cpu_bytes is not None:
        for key, d in (("cfg2_serialize", "serialize_gbs"), ("cfg4_deserialize", "deserialize_gbs")):
            if key in also and "error" not in cpu_bytes and "error" not in also[key]:
                also[key]["cpu_baseline"] = {
                    "value": cpu_bytes["all_cores"][d], "unit": "GB/s of serialized bytes", "cores": cpu_bytes["cores"],
                    "kind": cpu_bytes["kind"], "one_thread_gbs": cpu_bytes["one_thread"][d],
                    "sample": cpu_bytes["sample"] + " (the reference's Python loop, utils/__init__.py:193-276)"}
    if rank == 0:
        algo = CFG3_ALGO_BYTES_PER_IMAGE * n
        kern_s = t_local / args.steps
        line = {
            "metric": "images/s", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "cfg3: %d x 1920x1080 RGB u8 per GPU -> 224x224 fp32 NCHW VGG into a cudashm region" % n,
                       "global_batch": world * n, "parallelism": "shard-by-image x%d, one region per rank, no collective" % world,
                       "l2": "input 1.59 GB per step >> 126 MB L2 (no flush needed)",
                       "path": "fused" if info["fused_images"] == n else ("two-pass" if info["fused_images"] == 0 else "mixed"),
                       "parity_spot_check_all_ranks": parity},
            "roofline": roofline(algo, kern_s, peak, "k_preprocess_fused_cfg3", peak_source=peak_src,
                                 kernel="all kernels of one step (%d launch(es)/step)" % (launches // max(args.steps, 1))),
            "cpu_baseline": cpu,
            "e2e": None if args.no_e2e else {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": world * in_bytes,
                    "d2h_bytes_per_step": world * out_bytes, "steps": e2e_steps,
                    "path": "HostBatchPipeline.run: pinned host batch -> 8 chunks x (TcbMemcpy H2D | TcbPreprocessPlanRun "
                            "into the region | TcbMemcpy D2H of the result) on two streams",
                    "matches_device_run_all_ranks": e2e_ok,
                    "h2d_ceiling_gbs": h2d_ceiling_gbs,
                    "h2d_ceiling": "bare pinned->device cudaMemcpyAsync of the same %d bytes on all %d ranks at once" % (in_bytes, world),
                    "h2d_achieved_gbs": e2e_value * H * W * 3 / 1e9,
                    "frac_of_h2d_ceiling": e2e_value * H * W * 3 / 1e9 / h2d_ceiling_gbs,
                    "bidir_ceiling_gbs": bidir_ceiling_gbs,
                    "bidir_ceiling": "the same copy with the %d-byte result read back on a second stream at the same time "
                                     "(H2D GB/s until both are done)" % out_bytes,
                    "frac_of_bidir_ceiling": e2e_value * H * W * 3 / 1e9 / bidir_ceiling_gbs,
                    "affinity": affinity, "from_pageable": pageable},
            "gpu_launches": launches, "clocks": clocks, "also": also,
            "wall_s": time.perf_counter() - t_start,
        }
        print(json.dumps(line))
    plan.close()
    cudashm.destroy_shared_memory_region(region)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
