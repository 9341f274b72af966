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
    return torch.cat(out0)[:count].contiguous(), torch.cat(out1)[:count].contiguous()


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
        # NCCL prints its version banner on stdout at the VERSION level: keep stdout to the one JSON line
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    steps, warmup = args.steps, max(3, args.warmup)
    sp_path, lg_path = weight_files()
    pipe = get_model("two_view_pipeline")({
        "extractor": {**EXTRACTOR_CONF, "weights": sp_path},
        "matcher": {**MATCHER_CONF, "weights": lg_path},
    }).to(dev).eval()

    im0, im1 = make_pairs(rank * PAIRS, PAIRS)
    host = {"image0": im0.pin_memory(), "image1": im1.pin_memory(),
            "size": torch.tensor([[float(W), float(H)]] * PAIRS).pin_memory()}
    h2d_bytes = sum(t.numel() * t.element_size() for t in host.values())

    def to_views(i0, i1, size):
        return {"view0": {"image": i0, "image_size": size}, "view1": {"image": i1, "image_size": size}}

    resident = to_views(host["image0"].to(dev), host["image1"].to(dev), host["size"].to(dev))
    out_host = {}

    def step_resident():
        with torch.no_grad():
            return pipe(resident)

    # End to end: every step copies its own inputs from pinned host memory and returns its result record to
    # the host.  As a data loader would, the copy of step i+1 is issued on a side stream while step i computes
    # (two device input buffers); every copy and every read-back stays inside the timed region.
    copy_stream = torch.cuda.Stream(device=dev)
    dev_in = [{k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in host.items()} for _ in range(2)]
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    e2e_state = {"i": 0, "n": 0}

    def issue_copy(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])  # the step that last read this buffer has finished with it
            for k, v in host.items():
                dev_in[slot][k].copy_(v, non_blocking=True)
            copied[slot].record(copy_stream)

    def e2e_begin(n):
        e2e_state["i"], e2e_state["n"] = 0, n
        for ev in consumed:
            ev.record()

    def step_e2e():
        i, n = e2e_state["i"], e2e_state["n"]
        slot = i & 1
        if i == 0:
            issue_copy(0)
        if i + 1 < n:
            issue_copy(slot ^ 1)
        main = torch.cuda.current_stream()
        main.wait_event(copied[slot])
        with torch.no_grad():
            d = dev_in[slot]
            pred = pipe(to_views(d["image0"], d["image1"], d["size"]))
        consumed[slot].record(main)
        for k in RESULT_KEYS:
            if k not in out_host:
                out_host[k] = torch.empty(pred[k].shape, dtype=pred[k].dtype).pin_memory()
            out_host[k].copy_(pred[k], non_blocking=True)
        done = torch.cuda.Event()
        done.record(main)
        done.synchronize()  # the host reads the result record of this step (the prefetch keeps running)
        e2e_state["i"] = i + 1
        return pred

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
    e2e_begin(2)
    for _ in range(2):
        step_e2e()
    torch.cuda.synchronize()
    e2e_begin(steps)
    ms_e2e = timed(step_e2e, steps, profile=False)
    d2h_bytes = sum(t.numel() * t.element_size() for t in out_host.values())

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
                       "l2": "inputs_exceed_l2 (268 MB of images, >10 GB of activations per step)"},
            "e2e": {"value": total_pairs / (ms_e2e * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": ms_e2e / steps,
                    "pipelining": "step i+1's pinned-host -> device copy runs on a side stream under step i's kernels; "
                                  "all copies and read-backs are inside the timed region"},
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
            del dev_in
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the torch-eager GPU-library baseline")
    ap.add_argument("--cpu-batch", type=int, default=1, help="--impl reference: pairs per CPU step")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
