#!/usr/bin/env python
"""bench.py - throughput of the tensor-marshalling hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cfg3]

One JSON line on stdout (rank 0). A "step" is one pass of the hot path over one batch of
synthetic input. Default workload = cfg3 (BASELINE.json configs[2], the configuration the
images/s metric is quoted on): 256 synthetic 1920x1080 RGB uint8 images per GPU -> 224x224 fp32
NCHW, VGG mean subtraction, written into a CUDA shared-memory region. The BYTES configs
(cfg2 serialize 1M strings, cfg4 deserialize 16M strings) are measured in the same run and
reported under "also" with their own roofline objects.

Multi-GPU: launched under torchrun (one rank per GPU); the batch shards by image, every rank
fills its own region, there is no data-path collective (weak scaling). NCCL is used only for
the barrier and the max-over-ranks of the device-timed region.
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


# ---- clocks sampling ------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

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
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- CPU baseline: the reference's preprocess body (PIL + numpy) on the host cores ------------------
def _cpu_pre_worker(args):
    seed, count = args
    from PIL import Image

    from oracle import pil_bilinear as pb

    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (CFG3["H"], CFG3["W"], 3), dtype=np.uint8)
    img = Image.fromarray(a, "RGB")
    t0 = time.perf_counter()
    for _ in range(count):
        out = pb.preprocess_pil(img, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"], CFG3["scaling"])
        np.ascontiguousarray(out)
    return time.perf_counter() - t0


def cpu_preprocess_rate(images_per_worker, workers):
    """images/s of oracle.pil_bilinear.preprocess_pil (the reference's function body with the
    installed Pillow) over `workers` processes."""
    import multiprocessing as mp

    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        pool.map(_cpu_pre_worker, [(i, 1) for i in range(workers)])  # warm-up: imports, page-in
        t0 = time.perf_counter()
        pool.map(_cpu_pre_worker, [(1000 + i, images_per_worker) for i in range(workers)])
        dt = time.perf_counter() - t0
    return workers * images_per_worker / dt, dt


def cpu_images_per_worker(workers):
    """Bounded sample: about 15 s of CPU work per step (a 1080p image costs ~11-17 ms per core)."""
    return max(4, int(15.0 / 0.013 / max(workers, 1)))


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (PIL + numpy, restated
    in oracle/ because /root/reference does not travel to the GPU box) on all host cores."""
    if rank != 0:
        return
    cores = usable_cores()
    workers = max(1, cores)
    per = cpu_images_per_worker(workers)
    total_steps = args.warmup + args.steps
    rates = []
    t_all = time.perf_counter()
    for s in range(total_steps):
        rate, dt = cpu_preprocess_rate(per, workers)
        if s >= args.warmup:
            rates.append((rate, dt))
        if time.perf_counter() - t_all > 240:
            break
    value = float(np.mean([r for r, _ in rates])) if rates else rate
    ms = float(np.mean([d for _, d in rates])) * 1e3 if rates else dt * 1e3
    import PIL

    sample = "%d procs x %d images of 1920x1080 per step (PIL %s resize + numpy), %d timed steps" % (
        workers, per, PIL.__version__, len(rates))
    line = {
        "impl": "reference", "metric": "images/s", "value": value, "unit": "images/s", "n_gpus": args.gpus,
        "steps": len(rates), "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cfg3: 1920x1080 RGB u8 -> 224x224 fp32 NCHW VGG (bounded sample on host cores)"},
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": workers, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ---- BYTES side benchmarks (reported under "also") -------------------------------------------------
def bench_bytes(torch, dev, peak, steps, warmup):
    from triton_client_b200 import _lib
    from triton_client_b200 import utils as u

    res = {}
    # cfg2: serialize 1M strings (0..256 B), G-BYTES recipe, device-resident blob + int32 lengths
    rng = np.random.default_rng(42)
    lens = rng.integers(0, 257, 1_000_000, dtype=np.int64)
    blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
    d_blob = torch.from_numpy(blob).to("cuda:%d" % dev)
    d_lens = torch.from_numpy(lens.astype(np.int32)).to("cuda:%d" % dev)
    P, N = int(blob.size), int(lens.size)
    out = torch.empty(P + 4 * N, dtype=torch.uint8, device="cuda:%d" % dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:%d" % dev)
    times = []
    for i in range(warmup + steps):
        flush.fill_(i & 0xFF)  # L2 flush between iterations (264 MB working set ~ 2x L2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        u.serialize_to_device(d_blob, d_lens, out, 0, dev, sync=False)
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            times.append(e0.elapsed_time(e1) * 1e-3)
    t = float(np.median(times))
    algo = 2 * (P + 4 * N)
    res["cfg2_serialize"] = {
        "metric": "BYTES serialize GB/s", "value": (P + 4 * N) / t / 1e9, "unit": "GB/s of serialized output",
        "us_per_launch": t * 1e6, "elements": N,
        "roofline": {"bound": "hbm", "achieved": algo / t / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": algo / t / 1e9 / peak, "traffic": ncu_traffic("k_bytes_serialize_cfg2")},
        "config": {"workload": "cfg2: 1M strings len~U[0,256] (G-BYTES seed 42), device blob+int32 lengths -> device buffer",
                   "l2": "256 MiB flush write between iterations", "includes": "workspace memset + scan+scatter kernel"}}
    # cfg4: deserialize 16M strings (~2.1 GB); stream generated on the device by the serializer
    g = torch.Generator(device="cuda:%d" % dev).manual_seed(43)
    N4 = 16_000_000
    lens4 = torch.randint(0, 257, (N4,), dtype=torch.int32, device="cuda:%d" % dev, generator=g)
    P4 = int(lens4.sum(dtype=torch.int64).item())
    blob4 = torch.randint(0, 256, (P4,), dtype=torch.uint8, device="cuda:%d" % dev, generator=g)
    stream4, tot4 = u.serialize_to_device(blob4, lens4)
    del flush
    times = []
    launches0 = _lib.kernel_launch_count()
    for i in range(max(1, warmup // 2) + max(2, steps // 2)):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        offsets, payload = u.deserialize_from_device(stream4, tot4)
        e1.record()
        torch.cuda.synchronize()
        if i >= max(1, warmup // 2):
            times.append(e0.elapsed_time(e1) * 1e-3)
    ok = bool(offsets.numel() == N4 + 1 and payload.numel() == P4 and torch.equal(payload, blob4)
              and torch.equal(offsets[1:] - offsets[:-1], lens4.to(torch.int64)))
    t = float(np.median(times))
    algo4 = tot4 + P4 + 8 * (N4 + 1)
    res["cfg4_deserialize"] = {
        "metric": "BYTES deserialize GB/s", "value": tot4 / t / 1e9, "unit": "GB/s of serialized input",
        "ms_per_call": t * 1e3, "elements": N4, "roundtrip_bit_exact": ok,
        "roofline": {"bound": "hbm", "achieved": algo4 / t / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": algo4 / t / 1e9 / peak, "traffic": ncu_traffic("deserialize_cfg4_all_kernels")},
        "config": {"workload": "cfg4: 16M strings len~U[0,256] (~2.1 GB stream) -> int64 offsets + compact payload, all on device",
                   "l2": "input 2.1 GB >> L2", "includes": "all deserialize kernels + the 3 small D2H header reads the call makes"}}
    del stream4, blob4, lens4, offsets, payload
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--no-also", action="store_true", help="skip the BYTES side benchmarks")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3

    # CPU baseline first (rank 0, N=1), before CUDA is initialised in this process: the worker
    # pool is forked and only ever touches PIL / numpy.
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = usable_cores()
        per = cpu_images_per_worker(cores)
        rate, dt = cpu_preprocess_rate(per, cores)
        import PIL

        cpu = {"value": rate, "unit": "images/s", "cores": cores, "kind": "port",
               "sample": "%d procs x %d images of 1920x1080 (oracle.preprocess_pil = the reference's preprocess "
                         "body, Pillow %s + numpy %s), %.1f s" % (cores, per, PIL.__version__, np.__version__, dt)}

    import torch
    import torch.distributed as dist

    from triton_client_b200 import _lib
    from triton_client_b200.image_client import PreprocessPlan
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    dev = local_rank
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % dev))

    def barrier():
        if world > 1:
            dist.barrier()

    peak, peak_src = measured_peak_hbm()
    n = CFG3["n"]
    H, W = CFG3["H"], CFG3["W"]
    g = torch.Generator(device="cuda:%d" % dev).manual_seed(1000 + rank)
    batch = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda:%d" % dev, generator=g)
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
    barrier()
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()
    launches0 = _lib.kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    launches = _lib.kernel_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t_local = e0.elapsed_time(e1) * 1e-3
    tt = torch.tensor([t_local], dtype=torch.float64, device="cuda:%d" % dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_total = float(tt.item())
    ms_per_step = t_total / args.steps * 1e3
    value = world * n * args.steps / t_total
    info = plan.info()

    # parity spot check inside the bench (oracle as the checker only)
    parity = None
    if rank == 0:
        from oracle import clib
        from oracle import pil_bilinear as pb

        got = cudashm.get_contents_as_numpy(region, np.float32, [n, 3, 224, 224])[5]
        u8 = clib.resize_bilinear_u8(batch[5].cpu().numpy(), 224, 224)
        lut = pb.scaling_lut("FP32", 3, "VGG")
        parity = bool(np.array_equal(got, np.stack([lut[ch][u8[:, :, ch]] for ch in range(3)])))

    # ---- end to end: host buffers in, result read back, through the public API -------------------
    h_batch = torch.empty((n, H, W, 3), dtype=torch.uint8, pin_memory=True)
    h_batch.copy_(batch)
    h_out = torch.empty(out_bytes, dtype=torch.uint8, pin_memory=True)
    from triton_client_b200.image_client import HostBatchPipeline

    pipe = HostBatchPipeline(n, W, H, 3, CFG3["format"], CFG3["dtype"], CFG3["c"], CFG3["h"], CFG3["w"],
                             CFG3["scaling"], device=dev, chunks=8)

    def e2e_step():
        pipe.run(h_batch.data_ptr(), region._base_addr, out_bytes, h_out.data_ptr())

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(3):
        e2e_step()
    torch.cuda.synchronize()
    barrier()
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    te = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device="cuda:%d" % dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * n * e2e_steps / float(te.item())
    e2e_ok = None
    if rank == 0:  # the bytes that came back over PCIe are the ones the device-resident run produced
        ref_out = torch.empty(out_bytes, dtype=torch.uint8, pin_memory=True)
        step()
        _lib.check(_lib.lib.TcbRegionGet(region._c_handle, 0, ref_out.data_ptr(), out_bytes,
                                         torch.cuda.current_stream(dev).cuda_stream))
        e2e_ok = bool(torch.equal(ref_out, h_out))
    pipe.close()
    del h_batch, pipe

    also = {}
    if rank == 0 and world == 1:
        if not args.no_also:
            try:
                del batch
                torch.cuda.empty_cache()
                also = bench_bytes(torch, dev, peak, min(args.steps, 20), min(args.warmup, 5))
            except Exception as ex:  # the primary line must still be printed
                also = {"error": repr(ex)}

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
                       "parity_spot_check": parity},
            "roofline": {"bound": "hbm", "achieved": algo / kern_s / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": algo / kern_s / 1e9 / peak, "traffic": ncu_traffic("k_preprocess_fused_cfg3"),
                         "peak_source": peak_src,
                         "kernel": "all kernels of one step (%d launch(es)/step)" % (launches // max(args.steps, 1)),
                         "algorithmic_bytes_per_launch": algo},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": world * n * H * W * 3,
                    "d2h_bytes_per_step": world * out_bytes, "steps": e2e_steps,
                    "path": "HostBatchPipeline.run: pinned host batch -> 8 chunks x (TcbMemcpy H2D | TcbPreprocessPlanRun "
                            "into the region | TcbMemcpy D2H of the result) on two streams",
                    "matches_device_run": e2e_ok},
            "gpu_launches": launches, "clocks": clocks, "also": also,
        }
        print(json.dumps(line))
    plan.close()
    cudashm.destroy_shared_memory_region(region)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
