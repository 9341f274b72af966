"""Benchmark of the SuperPoint-open + LightGlue forward pass (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W            # B200 kernels
    python bench.py --impl reference --gpus N --steps K ...  # CPU oracle port on the host cores

One step = one pass of the hot path over one batch of 32 synthetic 1024x1024 image pairs at 2048
keypoints (extractor on both views + matcher, through TwoViewPipeline.forward).  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from glue_factory_b200 import synth  # noqa: E402

METRIC = "SP+LightGlue pairs/s @2048 kpts"
H = W = 1024
KPTS = 2048
PAIRS = 32
EXTRACTOR_CONF = dict(name="extractors.superpoint_open", max_num_keypoints=KPTS, detection_threshold=-1,
                      nms_radius=3, force_num_keypoints=False)
MATCHER_CONF = dict(name="matchers.lightglue", filter_threshold=0.1, flash=False)
# algorithmic FLOPs per pair: 2 x 177.86 (SuperPoint-open) + 229.8 (LightGlue)
GFLOP_PER_PAIR = 585.5
DOMINANT_TAG = "attention"  # the kernel with the largest share of the step (profiles/, "breakdown" in the JSON line)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sustained=p["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


def make_pairs(first, count, seed=1234):
    """`count` synthetic pairs: 8 seeded base pairs, the rest are flips of them (cheap on the host)."""
    base = min(count, 8)
    im0, im1, _ = synth.synthetic_batch(first, base, H, W, seed)
    out0, out1 = [im0], [im1]
    flips = [(-1,), (-2,), (-1, -2)]
    i = 0
    while sum(x.shape[0] for x in out0) < count:
        out0.append(torch.flip(im0, flips[i % 3]))
        out1.append(torch.flip(im1, flips[i % 3]))
        i += 1
    # 8-bit at the source, like images read from disk: every arm (B200 resident / end to end, torch eager, CPU oracle)
    # sees float32(value / 255) of the same uint8 pixels (numpy_image_to_torch, utils/image.py:112-121)
    return to_unit(quantize(torch.cat(out0)[:count])), to_unit(quantize(torch.cat(out1)[:count]))


def quantize(im):
    return (im * 255.0).round().clamp(0, 255).to(torch.uint8).contiguous()


def to_unit(u8):
    return (u8.float() / 255.0).contiguous()


def weight_files():
    d = tempfile.mkdtemp(prefix="b200glue_w_")
    sp, lg = os.path.join(d, "sp.pth"), os.path.join(d, "lg.pth")
    torch.save(synth.superpoint_open_state_dict(), sp)
    torch.save(synth.lightglue_state_dict(), lg)
    return sp, lg


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 6 or not parts[0].isdigit():
                continue
            sm.append(int(parts[0]))
            mx = int(parts[1])
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def quiet_nccl():
    """NCCL writes its version banner to STDOUT at NCCL_DEBUG=VERSION and above (WARN included); stdout must carry
    the one JSON line only.  Unset means silent; an explicit VERSION / WARN request is downgraded."""
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
        os.environ["NCCL_DEBUG"] = "NONE"


def dist_env():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    return rank, world, int(os.environ.get("LOCAL_RANK", 0))


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/ is a torch-CPU fp32 restatement of the reference, pinned against
# reference-generated vectors; the reference itself cannot travel to the GPU box)
# ---------------------------------------------------------------------------------------------
SP_ORACLE_CONF = dict(max_num_keypoints=KPTS, detection_threshold=-1, nms_radius=3)


def oracle_batch_fn(device="cpu"):
    """The oracle port on `device` for a batch of pairs: SuperPoint-open on both views, then LightGlue
    (what TwoViewPipeline does in the reference, two_view_pipeline.py:72-90)."""
    from oracle import lightglue as olg
    from oracle import sp_open as osp

    sp_sd = {k: v.to(device) for k, v in synth.superpoint_open_state_dict().items()}
    lg_sd = {k: v.to(device) for k, v in synth.lightglue_state_dict().items()}

    def run(im0, im1):
        size = torch.tensor([[float(W), float(H)]] * im0.shape[0], device=device)
        with torch.no_grad():
            p0 = osp.forward(sp_sd, im0, SP_ORACLE_CONF)
            p1 = osp.forward(sp_sd, im1, SP_ORACLE_CONF)
            lg = olg.forward(lg_sd, p0["keypoints"], p1["keypoints"], p0["descriptors"], p1["descriptors"],
                             size, size, dict(filter_threshold=0.1))
        return p0, p1, lg

    return run


def time_cpu(num_pairs, warmup_pairs, batch=1, keep=0):
    """Times the oracle port on the host cores: `num_pairs` pairs in steps of `batch` pairs.
    Returns (pairs/s, seconds, threads, [(p0, p1, lg) of the first `keep` timed steps])."""
    torch.set_num_threads(os.cpu_count() or 1)
    run = oracle_batch_fn("cpu")
    im0, im1 = make_pairs(0, min(8, max(batch, num_pairs)))
    n = im0.shape[0]

    def chunk(i):
        idx = [(i * batch + j) % n for j in range(batch)]
        return im0[idx], im1[idx]

    for i in range(-(-warmup_pairs // batch)):
        run(*chunk(i))
    kept = []
    steps = max(1, num_pairs // batch)
    t0 = time.perf_counter()
    for i in range(steps):
        out = run(*chunk(i))
        if i < keep:
            kept.append(out)
    dt = time.perf_counter() - t0
    return steps * batch / dt, dt, torch.get_num_threads(), kept


def time_eager_gpu(dev, im0, im1, steps, warmup, chunk=8):
    """GPU-library baseline: the same reference algorithm through PyTorch eager on this
    B200 (cuDNN / cuBLAS / ATen kernels), i.e. the oracle port fed CUDA tensors.  One step = the same 32 resident
    pairs as the B200 arm, run in chunks of `chunk` pairs (the fp32 full-resolution activations of 64 images
    would need > 50 GB).  Timed with CUDA events; returns one record per numeric mode."""
    run = oracle_batch_fn(dev)
    out = []
    P = im0.shape[0]

    def step():
        for c in range(0, P, chunk):
            run(im0[c:c + chunk], im1[c:c + chunk])

    for name, tf32, autocast in (("fp32", False, False), ("tf32", True, False), ("bf16-autocast", True, True)):
        old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        try:
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                for _ in range(warmup):
                    step()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(steps):
                    step()
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            out.append({"dtype": name, "value": P / (ms * 1e-3), "unit": "pairs/s", "ms_per_step": ms})
        except Exception as e:  # a library failure must not lose the B200 line
            out.append({"dtype": name, "error": f"{type(e).__name__}: {e}"[:200]})
        finally:
            torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
        torch.cuda.empty_cache()
    return out


def _pct(x, q):
    x = x.flatten().float()
    k = max(1, min(x.numel(), int(round(q * x.numel()))))
    return x.kthvalue(k).values.item()


def parity_report(pipe, dev, im0, im1, ref):
    """Achieved parity of the B200 path against the oracle on the pairs the CPU leg just ran (same synthetic
    pairs, same weights, configs[1] sizes).  `ref` = (p0, p1, lg) oracle predictions of those pairs."""
    p0, p1, lg = ref
    B = im0.shape[0]
    size = torch.tensor([[float(W), float(H)]] * B, device=dev)
    views = {"view0": {"image": im0.to(dev), "image_size": size}, "view1": {"image": im1.to(dev), "image_size": size}}
    with torch.no_grad():
        pred = pipe(views)
        matcher = pipe.matcher
        tf = matcher({"keypoints0": p0["keypoints"].to(dev), "keypoints1": p1["keypoints"].to(dev),
                      "descriptors0": p0["descriptors"].to(dev), "descriptors1": p1["descriptors"].to(dev),
                      "view0": {"image_size": size}, "view1": {"image_size": size}})
        fin = matcher.assign_from_descriptors(lg["ref_descriptors0"][:, 0].to(dev), lg["ref_descriptors1"][:, 0].to(dev))
    out = {"pairs": B, "against": "oracle port (fp32, CPU) on the same pairs and weights"}
    heat = pred["heatmap0"].cpu()
    out["heatmap_max_abs_err_over_max"] = ((heat - p0["heatmap"]).abs().max() / p0["heatmap"].max()).item()
    overlap, e2e_hits, e2e_total = [], 0, 0
    for b in range(B):
        ours0 = {tuple(k): i for i, k in enumerate(pred["keypoints0"][b].cpu().tolist())}
        ours1 = [tuple(k) for k in pred["keypoints1"][b].cpu().tolist()]
        ref0 = [tuple(k) for k in p0["keypoints"][b].tolist()]
        ref1 = [tuple(k) for k in p1["keypoints"][b].tolist()]
        overlap.append(len(set(ours0) & set(ref0)) / len(ref0))
        m_ours, m_ref = pred["matches0"][b].cpu().tolist(), lg["matches0"][b].tolist()
        for i, j in enumerate(m_ref):  # end to end: the reference's matches, found again by coordinates
            if j < 0:
                continue
            e2e_total += 1
            k = ours0.get(ref0[i])
            if k is not None and m_ours[k] >= 0 and ours1[m_ours[k]] == ref1[j]:
                e2e_hits += 1
    out["keypoint_set_overlap"] = sum(overlap) / B
    out["e2e_reference_matches_reproduced"] = e2e_hits / max(1, e2e_total)
    out["reference_matches_per_pair"] = e2e_total / B
    # matcher teacher-forced on the oracle's keypoints / descriptors (bf16 operands through 9 layers)
    la, la_ref = tf["log_assignment"].cpu(), lg["log_assignment"]
    rel = (la - la_ref).abs() / la_ref.abs().clamp_min(1e-3)
    out["teacher_forced_matcher"] = {
        "match_agreement": (tf["matches0"].cpu() == lg["matches0"]).float().mean().item(),
        "log_assignment_rel_err": {"median": rel.median().item(), "p99": _pct(rel, 0.99), "max": rel.max().item()},
        "ref_descriptors_rel_l2": ((tf["ref_descriptors0"].cpu() - lg["ref_descriptors0"]).norm()
                                   / lg["ref_descriptors0"].norm()).item(),
    }
    # final stage teacher-forced on the oracle's last-layer descriptors (fp32-grade split GEMMs): 1e-3 target
    rel = (fin[0].cpu() - la_ref).abs() / la_ref.abs().clamp_min(1e-3)
    out["teacher_forced_assignment"] = {
        "match_agreement": (fin[1].cpu() == lg["matches0"]).float().mean().item(),
        "log_assignment_rel_err": {"median": rel.median().item(), "p99": _pct(rel, 0.99), "max": rel.max().item()},
    }
    return out


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    steps, warmup = args.steps, args.warmup
    batch = max(1, args.cpu_batch)
    pps, dt, cores, _ = time_cpu(steps * batch, warmup * batch, batch=batch)
    sample = (f"{steps} timed step(s) of {batch} pair(s) each (1024x1024, 2048 kpts, fp32) after {warmup} warm-up "
              f"step(s); one process with {cores} torch threads whatever --gpus says")
    line = {
        "impl": "reference", "metric": METRIC, "value": pps, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1000.0 * dt / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: SP-open + LightGlue, 2048 kpts, 1024x1024 synthetic pairs; "
                               f"CPU arm runs {batch} pair(s) per step", "pairs_per_step": batch},
        "cpu_baseline": {"value": pps, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": pps, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))



class E2EStepper:
    """End to end: every step copies its own inputs from pinned host memory and returns its result record to the
    host.  As a data loader would, the copy of step i+1 is issued on a side stream while step i computes (two
    device input buffers); every copy and every read-back stays inside the timed region."""

    def __init__(self, dev, host_batches, run, result_keys):
        """host_batches: list of dicts of pinned CPU tensors (cycled); run(dict of device tensors) -> pred."""
        self.dev, self.host, self.run, self.keys = dev, host_batches, run, result_keys
        self.copy_stream = torch.cuda.Stream(device=dev)
        first = host_batches[0]
        self.dev_in = [{k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in first.items()} for _ in range(2)]
        self.copied = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.done = [torch.cuda.Event() for _ in range(2)]
        self.out_host = {}
        self.i = self.n = 0
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in first.values())

    def _issue_copy(self, slot, step):
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])  # the step that last read this buffer has finished with it
            for k, v in self.host[step % len(self.host)].items():
                self.dev_in[slot][k].copy_(v, non_blocking=True)
            self.copied[slot].record(self.copy_stream)

    def begin(self, n):
        self.i, self.n = 0, n
        for ev in self.consumed:
            ev.record()

    def step(self):
        i, n = self.i, self.n
        slot = i & 1
        if i == 0:
            self._issue_copy(0, 0)
        if i + 1 < n:
            self._issue_copy(slot ^ 1, i + 1)
        main = torch.cuda.current_stream()
        main.wait_event(self.copied[slot])
        with torch.no_grad():
            pred = self.run(self.dev_in[slot])
        self.consumed[slot].record(main)
        out = self.out_host.setdefault(slot, {})
        for k in self.keys:
            if k not in out or out[k].shape != pred[k].shape:
                out[k] = torch.empty(pred[k].shape, dtype=pred[k].dtype).pin_memory()
            out[k].copy_(pred[k], non_blocking=True)
        self.done[slot].record(main)
        # the host reads the result record of the PREVIOUS step while this one computes (two pinned result buffers),
        # so the launch queue never runs dry; the last step's record is read before the timed region ends
        if i > 0:
            self.done[slot ^ 1].synchronize()
        if i + 1 == n:
            self.done[slot].synchronize()
        self.i = i + 1
        return pred

    def d2h_bytes(self):
        return sum(t.numel() * t.element_size() for t in self.out_host.get(0, {}).values())


def make_timer(dev, world):
    import torch.distributed as dist

    from glue_factory_b200 import ops

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n, profile, only=None):
        barrier()
        ops.reset_stats(profile=profile, only=only)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    return barrier, timed


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
RESULT_KEYS = ["keypoints0", "keypoints1", "keypoint_scores0", "keypoint_scores1", "matches0", "matches1",
               "matching_scores0", "matching_scores1"]


def run_b200(args):
    import torch.distributed as dist

    from glue_factory_b200 import ops
    from gluefactory.models import get_model

    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        quiet_nccl()
        dist.init_process_group("nccl", device_id=dev)
    steps, warmup = args.steps, max(3, args.warmup)
    sp_path, lg_path = weight_files()
    pipe = get_model("two_view_pipeline")({
        "extractor": {**EXTRACTOR_CONF, "weights": sp_path},
        "matcher": {**MATCHER_CONF, "weights": lg_path},
    }).to(dev).eval()

    im0, im1 = make_pairs(rank * PAIRS, PAIRS)
    # end to end ships the uint8 pixels (a quarter of the float bytes); the first kernel converts them in its load
    host = {"image0": quantize(im0).pin_memory(), "image1": quantize(im1).pin_memory(),
            "size": torch.tensor([[float(W), float(H)]] * PAIRS).pin_memory()}

    def to_views(i0, i1, size):
        return {"view0": {"image": i0, "image_size": size}, "view1": {"image": i1, "image_size": size}}

    resident = to_views(im0.to(dev), im1.to(dev), host["size"].to(dev))  # float32 [0, 1], the forward(data) contract

    def step_resident():
        with torch.no_grad():
            return pipe(resident)

    e2e = E2EStepper(dev, [host], lambda d: pipe(to_views(d["image0"], d["image1"], d["size"])), RESULT_KEYS)
    barrier, timed = make_timer(dev, world)

    for _ in range(warmup):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    # timed region: CUDA events bracket the dominant kernel only (roofline); the per-kernel breakdown of every
    # launch comes from a second, untimed pass (two event records per launch lengthen a 130-launch step by ~4 %)
    ms = timed(step_resident, steps, profile=True, only=[DOMINANT_TAG])
    clocks = sampler.stop()
    launches = ops.launches()
    top_records = list(ops.profile_records())
    ms_profiled = timed(step_resident, steps, profile=True)
    records = list(ops.profile_records())
    ops.reset_stats(profile=False)
    e2e.begin(2)
    for _ in range(2):
        e2e.step()
    torch.cuda.synchronize()
    e2e.begin(steps)
    ms_e2e = timed(e2e.step, steps, profile=False)
    d2h_bytes, h2d_bytes = e2e.d2h_bytes(), e2e.h2d_bytes

    # final result gather (outside the timed region): fixed-size records per pair
    gather_ms = None
    if world > 1:
        rec = step_resident()["matches0"].to(torch.int32).contiguous()
        bucket = torch.empty(world * rec.numel(), dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dist.all_gather_into_tensor(bucket, rec.view(-1))
        torch.cuda.synchronize()
        gather_ms = 1000.0 * (time.perf_counter() - t0)

    if rank == 0:
        pk = peaks()
        per_tag = {}
        for tag, work, a, b in records:
            t = per_tag.setdefault(tag, [0.0, 0, 0.0])
            t[0] += a.elapsed_time(b)
            t[1] += 1
            t[2] += work
        ranked = sorted(per_tag.items(), key=lambda kv: -kv[1][0])
        top_tag = DOMINANT_TAG  # checked against the full breakdown below
        top_ms = sum(a.elapsed_time(b) for _, _, a, b in top_records)
        top_n = len(top_records)
        top_work = sum(w for _, w, _, _ in top_records)
        tensor_bound = top_tag.startswith(("conv3x3", "gemm", "attention"))
        achieved = top_work / (top_ms * 1e-3) / (1e12 if tensor_bound else 1e9)
        peak = pk["tf_sustained"] if tensor_bound else pk["hbm"]
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get(top_tag)
        total_pairs = PAIRS * world * steps
        value = total_pairs / (ms * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "configs[1]: SuperPoint-open + LightGlue, 2048 kpts, 1024x1024 synthetic pairs, "
                                   "batch 32 pairs per GPU per step", "pairs_per_step_per_gpu": PAIRS,
                       "image": [H, W], "keypoints": KPTS, "extractor": {k: v for k, v in EXTRACTOR_CONF.items()},
                       "matcher": MATCHER_CONF, "parallelism": f"pairs sharded over {world} GPU(s), no collective in the step",
                       "l2": "inputs_exceed_l2 (268 MB of images, >10 GB of activations per step)",
                       "images": "8-bit at the source: resident arm float32(value/255), end-to-end arm ships the uint8 pixels"},
            "e2e": {"value": total_pairs / (ms_e2e * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": ms_e2e / steps,
                    "pipelining": "step i+1's pinned-host -> device copy (uint8 pixels) runs on a side stream under step i's "
                                  "kernels; the host reads step i-1's result record while step i computes; all copies and "
                                  "read-backs are inside the timed region"},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "tensor" if tensor_bound else "hbm", "kernel": top_tag, "achieved": achieved,
                         "peak": peak, "unit": "TFLOP/s" if tensor_bound else "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": pk["src"] + (" sustained bf16" if tensor_bound else " copy"),
                         "launches_in_timed_region": top_n, "avg_ms": top_ms / top_n, "share_of_step": top_ms / ms},
            "pipeline": {"tflops": GFLOP_PER_PAIR * value / world / 1e3, "frac_of_sustained_peak":
                         GFLOP_PER_PAIR * value / world / 1e3 / pk["tf_sustained"]},
            "kernels": [{"tag": t, "ms_per_step": v[0] / steps, "launches_per_step": v[1] / steps,
                         "rate": v[2] / (v[0] * 1e-3) / 1e12 if v[0] > 0 else None} for t, v in ranked[:16]],
            "breakdown": {"how": "second pass of the same K steps with CUDA events around every launch, outside the "
                                 "timed region", "ms_per_step": ms_profiled / steps, "dominant": ranked[0][0]},
            "gather_ms": gather_ms,
        }
        if world == 1 and not args.no_gpu_baseline:
            e2e.dev_in = None
            torch.cuda.empty_cache()
            line["gpu_library_baseline"] = {
                "kind": "torch-eager", "what": "the oracle port (= the reference algorithm, plain torch ops) on CUDA "
                "tensors of this B200: cuDNN / cuBLAS / ATen kernels, same 32 resident pairs per step in chunks of 8",
                "steps": 3, "warmup": 3, "modes": time_eager_gpu(dev, resident["view0"]["image"],
                                                                 resident["view1"]["image"], 3, 3)}
        if world == 1 and not args.no_cpu_baseline:
            pps, dt, cores, kept = time_cpu(10, 1, batch=1, keep=2)
            line["cpu_baseline"] = {"value": pps, "unit": "pairs/s", "cores": cores, "kind": "port",
                                    "sample": "10 timed pairs (batch 1, as eval/hpatches.py:32 runs it) after 1 warm-up pair "
                                              "of the same workload (1024x1024, 2048 kpts), oracle port, fp32"}
            pps8, dt8, _, _ = time_cpu(8, 0, batch=8)
            line["cpu_baseline"]["batch8"] = {"value": pps8, "unit": "pairs/s", "sample": "1 timed step of 8 pairs"}
            # parity of the B200 arm on the two pairs the CPU leg just computed (oracle = checker only)
            cat = lambda i, key: torch.cat([k[i][key] for k in kept])  # noqa: E731
            ref = tuple({key: cat(i, key) for key in kept[0][i] if torch.is_tensor(kept[0][i][key])} for i in range(3))
            pim0, pim1 = make_pairs(0, 2)
            line["parity"] = parity_report(pipe, dev, pim0, pim1, ref)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# other BASELINE.json configurations: --workload adaptive (configs[2]), superglue (configs[3]), sweep (configs[4])
# ---------------------------------------------------------------------------------------------
def _hash_uniform(ids, count, salt, dev):
    """Stateless uniform numbers in [0, 1): element e of pair id i depends only on (i, salt, e), so a pair's image is
    the same whichever rank / batch generates it.  ids: int64 [P] -> float32 [P, count]."""
    e = torch.arange(count, dtype=torch.int64, device=dev)[None]
    v = (ids[:, None] * 1000003 + salt * 7919 + e * 2654435761) & 0xFFFFFFFF
    for mul in (0x7FEB352D, 0x846CA68B):  # two rounds of a 32-bit integer mix
        v = ((v ^ (v >> 16)) * mul) & 0xFFFFFFFF
    v = v ^ (v >> 16)
    return v.float() / 4294967296.0


def device_pairs(ids, dev):
    """Distinct synthetic pairs generated on the device (the sweep needs 8192 of them; the host recipe of
    synth.synthetic_pair takes ~0.2 s per pair): band-limited noise octaves, a seeded homography warp and noise, as
    SURVEY 8d describes, quantised to uint8.  ids: list of global pair ids -> two uint8 tensors [P,1,H,W]."""
    import torch.nn.functional as F

    ids_t = torch.tensor(ids, dtype=torch.int64, device=dev)
    P = len(ids)
    acc = torch.zeros(P, 1, H, W, device=dev)
    for salt, (weight, div) in enumerate(((8.0, 32), (4.0, 16), (2.0, 8), (1.0, 4))):
        h, w = H // div, W // div
        octave = (_hash_uniform(ids_t, h * w, salt, dev).view(P, 1, h, w) - 0.5) * (12.0 ** 0.5)
        acc += weight * F.interpolate(octave, size=(H, W), mode="bilinear", align_corners=False)
    lo, hi = acc.amin((1, 2, 3), keepdim=True), acc.amax((1, 2, 3), keepdim=True)
    im0 = (acc - lo) / (hi - lo)
    r = _hash_uniform(ids_t, 12, 17, dev)  # rotation, zoom 0.95, corner jitter folded into an affine + projective part
    ang = (r[:, 0] - 0.5) * 0.1
    a = torch.zeros(P, 3, 3, device=dev)
    a[:, 0, 0], a[:, 0, 1], a[:, 0, 2] = torch.cos(ang) / 0.95, -torch.sin(ang) / 0.95, (r[:, 1] - 0.5) * 0.04
    a[:, 1, 0], a[:, 1, 1], a[:, 1, 2] = torch.sin(ang) / 0.95, torch.cos(ang) / 0.95, (r[:, 2] - 0.5) * 0.04
    a[:, 2, 0], a[:, 2, 1], a[:, 2, 2] = (r[:, 3] - 0.5) * 0.02, (r[:, 4] - 0.5) * 0.02, 1.0
    ys, xs = torch.meshgrid(torch.linspace(-1, 1, H, device=dev), torch.linspace(-1, 1, W, device=dev), indexing="ij")
    pts = torch.stack([xs, ys, torch.ones_like(xs)], -1).view(1, H * W, 3) @ a.transpose(1, 2)
    grid = (pts[..., :2] / pts[..., 2:]).view(P, H, W, 2)
    im1 = F.grid_sample(im0, grid, mode="bilinear", padding_mode="border", align_corners=False)
    im1 = (im1 + (_hash_uniform(ids_t, H * W, 29, dev).view(P, 1, H, W) - 0.5) * 0.02).clamp(0, 1)
    return quantize(im0), quantize(im1)


WORKLOADS = {
    "adaptive": dict(
        metric="SP+LightGlue adaptive (depth / width confidence 0.95) pairs/s @4096 kpts", pairs=8, kpts=4096, tag="attention",
        workload="configs[2]: SuperPoint-open + LightGlue with early exit and point pruning (depth_confidence 0.95, "
                 "width_confidence 0.95), 4096 kpts, 1024x1024 synthetic pairs, batch 8 pairs per GPU per step"),
    "superglue": dict(
        metric="SP+SuperGlue pairs/s @2048 kpts, 100 Sinkhorn iterations", pairs=32, kpts=2048, tag="sg_sinkhorn",
        workload="configs[3]: SuperPoint (gluefactory_nonfree) + SuperGlue, log-domain Sinkhorn, 100 iterations, "
                 "2048 kpts, 1024x1024 synthetic pairs, batch 32 pairs per GPU per step"),
    "sweep": dict(
        metric=METRIC, pairs=PAIRS, kpts=KPTS, tag="attention",
        workload="configs[4]: pair-sharded sweep, SuperPoint-open + LightGlue, 2048 kpts, over distinct synthetic pairs "
                 "(pair i -> rank i mod N), batches of 32 pairs"),
}


def build_pipe(name, dev):
    from gluefactory.models import get_model

    d = tempfile.mkdtemp(prefix="b200glue_w_")
    if name == "superglue":
        torch.save(synth.superpoint_nonfree_state_dict(), d + "/spn.pth")
        pipe = get_model("two_view_pipeline")({
            "extractor": {"name": "gluefactory_nonfree.superpoint", "max_num_keypoints": 2048, "nms_radius": 3,
                          "detection_threshold": -1, "weights": d + "/spn.pth"},
            "matcher": {"name": "gluefactory_nonfree.superglue", "weights": None, "num_sinkhorn_iterations": 100,
                        "filter_threshold": 0.2}})
        pipe.matcher.load_state_dict(synth.superglue_state_dict())
        return pipe.to(dev).eval()
    sp, lg = weight_files()
    ext = {**EXTRACTOR_CONF, "weights": sp, "max_num_keypoints": WORKLOADS[name]["kpts"]}
    mat = {**MATCHER_CONF, "weights": lg}
    if name == "adaptive":
        mat.update(depth_confidence=0.95, width_confidence=0.95)
    return get_model("two_view_pipeline")({"extractor": ext, "matcher": mat}).to(dev).eval()


def cpu_pairs_per_s(name, im0, im1, n):
    """The oracle port of the workload on the host cores, batch 1, `n` timed pairs after one warm-up pair."""
    from oracle import lightglue as olg
    from oracle import sp_nonfree as ospn
    from oracle import sp_open as osp
    from oracle import superglue as osg

    torch.set_num_threads(os.cpu_count() or 1)
    size = torch.tensor([[float(W), float(H)]])
    if name == "superglue":
        spn, sg = synth.superpoint_nonfree_state_dict(), synth.superglue_state_dict()
        conf = dict(max_num_keypoints=2048, nms_radius=3, detection_threshold=-1)

        def run(a, b):
            p0, p1 = ospn.forward(spn, a, conf), ospn.forward(spn, b, conf)
            return osg.forward(sg, p0["keypoints"], p1["keypoints"], p0["descriptors"], p1["descriptors"],
                               p0["keypoint_scores"], p1["keypoint_scores"], size, size,
                               dict(num_sinkhorn_iterations=100, filter_threshold=0.2))
    else:
        sp, lg = synth.superpoint_open_state_dict(), synth.lightglue_state_dict()
        conf = dict(max_num_keypoints=WORKLOADS[name]["kpts"], detection_threshold=-1, nms_radius=3)
        lconf = dict(filter_threshold=0.1)
        if name == "adaptive":
            lconf.update(depth_confidence=0.95, width_confidence=0.95)

        def run(a, b):
            p0, p1 = osp.forward(sp, a, conf), osp.forward(sp, b, conf)
            return olg.forward(lg, p0["keypoints"], p1["keypoints"], p0["descriptors"], p1["descriptors"], size, size, lconf)
    with torch.no_grad():
        run(im0[:1], im1[:1])
        t0 = time.perf_counter()
        for i in range(n):
            run(im0[i % len(im0)][None], im1[i % len(im0)][None])
        dt = time.perf_counter() - t0
    return n / dt, torch.get_num_threads()


def run_workload(args):
    import torch.distributed as dist

    from glue_factory_b200 import ops, shard
    from glue_factory_b200.models.lightglue_adaptive import executed_flops

    name = args.workload
    wl = WORKLOADS[name]
    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        quiet_nccl()
        dist.init_process_group("nccl", device_id=dev)
    P, K = wl["pairs"], wl["kpts"]
    steps, warmup = args.steps, max(3, args.warmup)
    pipe = build_pipe(name, dev)
    size = torch.tensor([[float(W), float(H)]] * P)
    barrier, timed = make_timer(dev, world)

    def to_views(i0, i1, sz):
        return {"view0": {"image": i0, "image_size": sz}, "view1": {"image": i1, "image_size": sz}}

    total_pairs, ids = None, None
    if name == "sweep":
        # distinct pairs, pair i -> rank i mod N (shard.pairs_for_rank); resident as uint8 in HBM (2 MB per pair)
        total_pairs = args.sweep_pairs
        ids = shard.pairs_for_rank(total_pairs, rank, world)
        ids = ids[: len(ids) // P * P]
        steps = len(ids) // P
        batches = [device_pairs(ids[i * P: (i + 1) * P], dev) for i in range(steps)]
        size_d = size.to(dev)
        keep_m = torch.empty(len(ids), K, dtype=torch.int64, device=dev)
        keep_s = torch.empty(len(ids), K, dtype=torch.float32, device=dev)
        cursor = {"i": 0}

        def step_resident():
            i = cursor["i"] % steps
            with torch.no_grad():
                pred = pipe(to_views(batches[i][0], batches[i][1], size_d))
            keep_m[i * P: (i + 1) * P] = pred["matches0"]
            keep_s[i * P: (i + 1) * P] = pred["matching_scores0"]
            cursor["i"] += 1
            return pred

        ring = min(steps, 16)  # pinned host ring of the first batches (2 MB per pair), cycled by the end-to-end arm
        host = [{"image0": batches[i][0].cpu().pin_memory(), "image1": batches[i][1].cpu().pin_memory(),
                 "size": size.pin_memory()} for i in range(ring)]
        im0 = im1 = None
    else:
        im0, im1 = make_pairs(rank * P, P)
        resident = to_views(im0.to(dev), im1.to(dev), size.to(dev))

        def step_resident():
            with torch.no_grad():
                return pipe(resident)

        host = [{"image0": quantize(im0).pin_memory(), "image1": quantize(im1).pin_memory(), "size": size.pin_memory()}]
    e2e = E2EStepper(dev, host, lambda d: pipe(to_views(d["image0"], d["image1"], d["size"])), RESULT_KEYS)

    for _ in range(warmup):
        step_resident()
    if name == "sweep":
        cursor["i"] = 0
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(step_resident, steps, profile=True, only=[wl["tag"]])
    clocks = sampler.stop()
    launches = ops.launches()
    top_records = list(ops.profile_records())
    stats = executed_flops(pipe.matcher.adaptive_stats) if name == "adaptive" else None
    if name == "sweep":
        cursor["i"] = 0
    ms_profiled = timed(step_resident, min(steps, 10), profile=True)
    records = list(ops.profile_records())
    ops.reset_stats(profile=False)
    e2e_steps = steps if name != "sweep" else min(steps, 32)
    e2e.begin(2)
    for _ in range(2):
        e2e.step()
    torch.cuda.synchronize()
    e2e.begin(e2e_steps)
    ms_e2e = timed(e2e.step, e2e_steps, profile=False)

    gather_ms, checksum = None, None
    if name == "sweep":
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        all_m, all_s = shard.gather_match_records(keep_m, keep_s, ids, total_pairs, world)  # the one collective: final gather
        torch.cuda.synchronize()
        gather_ms = 1000.0 * (time.perf_counter() - t0)
        covered = world * len(ids)
        # invariant under the number of ranks: every pair's matches depend on the pair id only
        checksum = {"pairs_gathered": int(covered), "matches": int((all_m >= 0).sum().item()),
                    "sum_match_index": int(all_m.clamp(min=0).sum().item()),
                    "sum_scores": float(all_s.double().sum().item())}

    if rank == 0:
        pk = peaks()
        per_tag = {}
        for tag, work, a, b in records:
            t = per_tag.setdefault(tag, [0.0, 0, 0.0])
            t[0] += a.elapsed_time(b)
            t[1] += 1
            t[2] += work
        n_prof = min(steps, 10)
        ranked = sorted(per_tag.items(), key=lambda kv: -kv[1][0])
        top_ms = sum(a.elapsed_time(b) for _, _, a, b in top_records)
        top_n = max(1, len(top_records))
        top_work = sum(w for _, w, _, _ in top_records)
        tensor_bound = wl["tag"] == "attention"
        if name == "adaptive":  # executed attention FLOPs: live lengths per layer, not the padded launch sizes
            hist = pipe.matcher.adaptive_stats["hist"].cpu().double()
            m, n = hist[:, :P], hist[:, P:]
            top_work = float((1024.0 * (m * m + n * n) + 1536.0 * m * n).sum()) * steps
        if name == "superglue":  # SURVEY 8d streaming model: 2 passes x iterations over the (M+1)(N+1) fp32 couplings
            top_work = 2.0 * 100 * (K + 1) * (K + 1) * 4.0 * P * steps
        achieved = top_work / (top_ms * 1e-3) / (1e12 if tensor_bound else 1e9)
        peak = pk["tf_sustained"] if tensor_bound else pk["hbm"]
        pairs_done = P * world * steps
        value = pairs_done / (ms * 1e-3)
        line = {
            "metric": wl["metric"], "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak" if name != "sweep" else "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl["workload"], "pairs_per_step_per_gpu": P, "image": [H, W], "keypoints": K,
                       "parallelism": f"pairs sharded over {world} GPU(s), no collective in the step",
                       "l2": "inputs_exceed_l2"},
            "e2e": {"value": P * world * e2e_steps / (ms_e2e * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": e2e.h2d_bytes,
                    "d2h_bytes_per_step": e2e.d2h_bytes(), "ms_per_step": ms_e2e / e2e_steps, "steps": e2e_steps,
                    "images": "uint8 pixels from pinned host memory, converted in the first kernel"},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"bound": "tensor" if tensor_bound else "hbm", "kernel": wl["tag"], "achieved": achieved, "peak": peak,
                         "unit": "TFLOP/s" if tensor_bound else "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": pk["src"], "launches_in_timed_region": len(top_records), "avg_ms": top_ms / top_n,
                         "share_of_step": top_ms / ms},
            "kernels": [{"tag": t, "ms_per_step": v[0] / n_prof, "launches_per_step": v[1] / n_prof} for t, v in ranked[:14]],
            "breakdown": {"ms_per_step": ms_profiled / n_prof, "dominant": ranked[0][0]},
            "gather_ms": gather_ms,
        }
        if stats is not None:
            line["adaptive"] = stats
        if checksum is not None:
            line["sweep"] = {"total_pairs": total_pairs, **checksum,
                             "e2e_note": f"end-to-end arm cycles a pinned host ring of {len(host)} distinct batches"}
        if world == 1 and not args.no_cpu_baseline:
            if im0 is None:
                im0, im1 = make_pairs(0, 2)
            n = 2 if name != "sweep" else 4
            pps, cores = cpu_pairs_per_s(name, im0, im1, n)
            line["cpu_baseline"] = {"value": pps, "unit": "pairs/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} timed pairs (batch 1) after 1 warm-up pair of the same workload, oracle port, fp32"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the torch-eager GPU-library baseline")
    ap.add_argument("--cpu-batch", type=int, default=1, help="--impl reference: pairs per CPU step")
    ap.add_argument("--workload", default="lightglue", choices=["lightglue"] + sorted(WORKLOADS),
                    help="lightglue = BASELINE configs[1] (the headline, default); adaptive / superglue / sweep = configs[2..4]")
    ap.add_argument("--sweep-pairs", type=int, default=8192)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload != "lightglue":
        run_workload(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
