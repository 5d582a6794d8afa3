#!/usr/bin/env python3
"""bench.py — ORB extract + stereo/temporal Hamming match throughput (BASELINE.json metric).

A "step" = one pass of the hot path over one batch of synthetic stereo pairs (default: KITTI-shaped
1241x376, 2000 features, stereo + temporal matching = BASELINE.json configs[1]).
  value : stereo pairs/s with the input images already resident in HBM (b2orb_stereo_frontend_device),
          timed with CUDA events on the handle's stream, max over ranks.
  e2e   : the same metric through the host-buffer C-ABI call (b2orb_stereo_frontend): pinned host images in,
          H2D + kernels + D2H of every result slab inside the timed region (two handles on two host threads
          so copies overlap kernels, the way the reference drives two extractor instances from two threads).
  roofline : dominant kernel, algorithmic bytes / live CUDA-event duration vs the measured HBM peak.
          By default the call returns what the reference keeps of a frame (_curPixelsL/R, _curDescriptorsL/R,
          match / circular lists, _curFeatureMask: src/feature-tracker.cpp:210-218, 311-321); `--e2e-out full` returns
          every keypoint and descriptor of both eyes as well (reported as e2e_full in the default run).
  roofline : dominant kernel, algorithmic bytes / live CUDA-event duration vs the measured HBM peak, plus the
          instruction-issue roofline (ncu instruction counts committed under profiles/).
  cpu_baseline : the reference's CPU implementation of the path on this box's host cores, bounded sample: extraction by
          the reference's own src/ORBextractor.cc (oracle/_ref, compiled against a mock OpenCV) when that library is in
          the snapshot ("kind": "reference"), else the oracle port; per-stage means and, when cv2 imports, the real
          OpenCV kernels for the stages OpenCV provides.
`--impl reference` times only that CPU implementation with all host threads, on the same config.
Multi-GPU (torchrun): every rank runs the same per-GPU workload on its own GPU (weak scaling, no collective on
the data path); only per-pair count summaries are gathered.
"""
import argparse
import importlib
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

METRIC = "orb_extract_stereo_match_pairs_per_s"
UNIT = "stereo pairs/s"


def algorithmic_bytes(geom_levels, nfeat, temporal):
    """B_pair = 2*B_img + B_match; returns (B_pair, per-kernel bytes per PAIR)."""
    p = [w * h for (w, h) in geom_levels]
    b_img = p[0] + 2 * sum(p[1:]) + sum(p) + nfeat * (961 + 1369) + nfeat * (32 + 28)
    b_match = 2 * nfeat * 32 + nfeat * 8
    per_kernel = {
        "resize": 2 * (sum(p[:-1]) + sum(p[1:])),          # read level l-1, write level l
        "fast": 2 * sum(p),                                # read every pyramid pixel once
        "blur": 2 * 2 * sum(p),                            # read + write every level
        "describe": 2 * nfeat * (961 + 1369 + 32 + 28),
        "match_stereo": b_match,
        "ingest": 2 * p[0] * (1 + 6 + 1),                  # remap: 4-tap gather ~ one read of the frame + 6 map bytes + 1 write per pixel
    }
    return 2 * b_img + b_match * (3 if temporal else 1), per_kernel


def level_sizes(w, h, nlevels=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(float(s) * float(np.float32(sf)))
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        sm = [int(r[0]) for r in self.rows if r and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def bind_near_gpu(gpu_index):
    """Multi-GPU runs: keep this rank's threads — and with them its pinned host buffers, which Linux places on the
    node of the allocating thread — on the CPUs next to its GPU, so the H2D / D2H DMA of the e2e leg does not cross the
    socket interconnect.  Returns the previous affinity (restored before the CPU baseline) or None."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(gpu_index)
        try:
            bus = "%08X:%02X:%02X.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            hnd = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            hnd = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        ncpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(hnd, (ncpu + 63) // 64)
        cpus = [i for i in range(ncpu) if (mask[i // 64] >> (i % 64)) & 1]
        prev = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in prev]
        if cpus and len(cpus) < len(prev):
            os.sched_setaffinity(0, cpus)
            return prev, len(cpus)
    except Exception:
        pass
    return None, 0


def cpu_runner(orc):
    """(run_pairs function, kind): the reference's own extractor when oracle/_ref travelled with the snapshot."""
    try:
        if orc.ref_available():
            return orc.ref_run_pairs, "reference"
    except Exception:
        pass
    return orc.run_pairs, "port"


CPU_NOTE = ("extraction = the reference's src/ORBextractor.cc built against a mock OpenCV whose image primitives (FAST, resize, "
            "GaussianBlur, fastAtan2) and the BFMatcher are scalar -O3 restatements without -march (no SIMD); the OpenCV the "
            "reference links is SIMD-optimised and faster on those stages (see cv2_primitives_ms and cv2_hybrid_lower_bound: about 2x overall)")


def cpu_baseline(synth, orc, shape, cfg, temporal, workers, target_s):
    """CPU implementation on host cores, bounded sample sized from a 1-round probe (about `target_s` seconds)."""
    run, kind = cpu_runner(orc)
    args = (synth.MATCH_RATIO, synth.MIN_MATCH_DIST, synth.MAX_VERTICAL_PIXEL_DIST)
    probe_n = max(workers, 2)
    ls, rs = synth.sequence(shape, probe_n, seed=1)
    t = time.perf_counter()
    run(cfg, ls, rs, *args, workers, 1, temporal)
    dt = time.perf_counter() - t
    n = int(min(max(probe_n, target_s / max(dt, 1e-3) * probe_n), 64 * workers))
    n = max(workers, n // workers * workers)
    ls, rs = synth.sequence(shape, n, seed=1)
    t = time.perf_counter()
    run(cfg, ls, rs, *args, workers, 1, temporal)
    dt = time.perf_counter() - t
    # per-stage means (method of benchmark/bundleAdjustment/bundleAdjustment.cpp:174-318: steady_clock around every stage,
    # mean over the iterations) from the instrumented port on a few pairs, one thread
    k = min(n, 6)
    orc.run_pairs(cfg, ls[:k], rs[:k], *args, 1, 1, temporal)
    stages = {s_: round(v / k * 1e3, 3) for s_, v in orc.last_stage_seconds().items()}
    return n / dt, n, dt, kind, stages


def cv2_primitive_ms(synth, shape, nfeat):
    """Real OpenCV kernels (the cv2 wheel, SIMD / IPP) for the stages OpenCV provides, per stereo pair, one thread:
    resize chain, GaussianBlur of every level, BFMatcher Hamming nfeat x nfeat.  None when cv2 does not import."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    w, h = synth.SHAPES[shape][:2]
    ls, rs = synth.sequence(shape, 1, seed=1)
    sizes = level_sizes(w, h)
    reps = 5
    t_res = t_blur = 0.0
    for _ in range(reps):
        for img in (ls[0], rs[0]):
            t = time.perf_counter()
            pyr = [img]
            for (lw, lh) in sizes[1:]:
                pyr.append(cv2.resize(pyr[-1], (lw, lh), interpolation=cv2.INTER_LINEAR))
            t_res += time.perf_counter() - t
            t = time.perf_counter()
            for lv in pyr:
                cv2.GaussianBlur(lv, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
            t_blur += time.perf_counter() - t
    rng = np.random.default_rng(0)
    q = rng.integers(0, 256, (nfeat, 32), dtype=np.uint8); tr = rng.integers(0, 256, (nfeat, 32), dtype=np.uint8)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    t = time.perf_counter()
    for _ in range(reps):
        bf.match(q, tr)
    t_bf = time.perf_counter() - t
    ini = synth.SHAPES[shape][3]
    fast = cv2.FastFeatureDetector_create(int(ini), True)
    pyrs = []
    for img in (ls[0], rs[0]):
        pyr = [img]
        for (lw, lh) in sizes[1:]:
            pyr.append(cv2.resize(pyr[-1], (lw, lh), interpolation=cv2.INTER_LINEAR))
        pyrs.append(pyr)
    t = time.perf_counter()
    for _ in range(reps):
        for pyr in pyrs:
            for lv in pyr:
                fast.detect(lv)    # ONE whole-level call (the reference makes ~30x30-px per-cell calls): a lower bound
    t_fast = time.perf_counter() - t
    return {"pyramid": round(t_res / reps * 1e3, 3), "blur": round(t_blur / reps * 1e3, 3),
            "match_stereo": round(t_bf / reps * 1e3, 3), "fast_all_levels_whole_image_calls": round(t_fast / reps * 1e3, 3),
            "opencv": cv2.__version__, "threads": 1}


def other_configs(fe, synth, device, peak, dbg, checker):
    """Bounded lines for the BASELINE.json configs the headline does not cover (a few steps each): batch throughput with
    inputs resident in HBM, single-pair latency through the host C-ABI call, parity of pair 0 against the oracle, and for
    the cfsd shape the paced trigger of config 3 (one frame every 5 ms = the 200 Hz IMU rate, config/cfsd.yml:76)."""
    import torch
    out = {}
    plan = {"euroc": (64, 6), "cfsd": (64, 6), "kitti_yml": (32, 6), "4k": (4, 4), "real_cfsd": (64, 6), "real_kitti": (32, 6)}
    real = {"real_cfsd": ("cfsd", [("cfsd360p_left", "cfsd360p_right"), ("cfsd_set_110_left", "cfsd_set_110_right")]),
            "real_kitti": ("kitti_yml", [("kitti0", "kitti2")])}
    for name, (npairs, steps) in plan.items():
        w, h, nfeat, ini, mn = synth.SHAPES[real[name][0] if name in real else name]
        cfg = (nfeat, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn)
        temporal = "kitti" in name
        if name in real:
            # REAL camera frames shipped with the reference (decoded once into tests/golden/real/*.npz): the synthetic texture
            # is far denser in corners than any real frame, so this is the line that says what real input statistics cost
            try:
                fr = [(np.load(os.path.join(ROOT, "tests", "golden", "real", a + ".npz"))["image"],
                       np.load(os.path.join(ROOT, "tests", "golden", "real", b_ + ".npz"))["image"]) for a, b_ in real[name][1]]
            except OSError:
                continue
            ls = np.stack([fr[i % len(fr)][0] for i in range(npairs)]); rs = np.stack([fr[i % len(fr)][1] for i in range(npairs)])
            assert ls.shape[1:] == (h, w)
        else:
            ls, rs = synth.sequence(name, npairs, seed=11)
        pitch = (w + 31) // 32 * 32
        d_l = torch.zeros((npairs, h, pitch), dtype=torch.uint8, device="cuda"); d_r = torch.zeros_like(d_l)
        d_l[:, :, :w] = torch.from_numpy(ls).cuda(); d_r[:, :, :w] = torch.from_numpy(rs).cuda()
        hd = fe.Handle(*cfg, w, h, max_images=2 * npairs, device=device)
        for k_, v_ in dbg:
            hd.debug_set(k_, v_)
        st = torch.cuda.ExternalStream(hd.stream, device=torch.device("cuda", device))
        run = lambda: hd.stereo_frontend_device(d_l.data_ptr(), d_r.data_ptr(), npairs, w, h, pitch, pitch * h, temporal)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(steps):
            run()
        e1.record(st)
        torch.cuda.synchronize()
        value = npairs * steps / (e0.elapsed_time(e1) * 1e-3)
        b_pair, _ = algorithmic_bytes(level_sizes(w, h), nfeat, temporal)
        res = {"workload": f"{name} {w}x{h}, {nfeat} feats", "pairs_per_step": npairs, "steps": steps, "value": round(value, 1),
               "unit": UNIT, "frac_of_hbm_peak": round(b_pair * value / 1e9 / peak, 4)}
        hd.profile_enable(True)                                  # per-kernel device time of this shape (CUDA events between launches)
        for _ in range(2):
            run()
        hd.sync()
        prof = hd.profile_read()
        hd.profile_enable(False)
        res["kernel_us_per_pair"] = {k: round(v[0] * 1e3 / (2 * npairs), 3) for k, v in prof.items() if v[1]}
        if checker:
            from oracle import oracle as orc
            o = hd.stereo_download(npairs, hd.alloc_stereo_out(npairs))
            ex = orc.OracleExtractor(*cfg)
            kl, dl = ex(ls[0]); kr, dr = ex(rs[0])
            nl, nr = int(o["n_l"][0]), int(o["n_r"][0])
            ti, di = orc.bf_match(dl, dr)
            res["bit_exact_vs_oracle_pair0"] = bool(
                nl == len(kl) and nr == len(kr) and o["kp_l"][0, :nl].tobytes() == kl.tobytes() and np.array_equal(o["desc_l"][0, :nl], dl)
                and o["kp_r"][0, :nr].tobytes() == kr.tobytes() and np.array_equal(o["desc_r"][0, :nr], dr)
                and np.array_equal(o["train_idx"][0, :nl], ti) and np.array_equal(o["dist"][0, :nl], di))
            # the reference's CPU implementation on the same frames: one round of min(pairs, host cores) pairs, one per thread
            run_cpu, kind_cpu = cpu_runner(orc)
            ncpu = min(npairs, os.cpu_count() or 1)
            t0 = time.perf_counter()
            run_cpu(cfg, ls[:ncpu], rs[:ncpu], synth.MATCH_RATIO, synth.MIN_MATCH_DIST, synth.MAX_VERTICAL_PIXEL_DIST, ncpu, 1, temporal)
            res["cpu_baseline"] = {"value": round(ncpu / (time.perf_counter() - t0), 2), "unit": UNIT, "cores": ncpu, "kind": kind_cpu,
                                   "sample": f"{ncpu} pairs, one per thread"}
        hd.close()
        # single-pair latency, pinned host images in, kept result set out
        hl = fe.Handle(*cfg, w, h, max_images=2, device=device)
        for k_, v_ in dbg:
            hl.debug_set(k_, v_)
        pl = fe.pinned_empty(ls.shape); pl[...] = ls
        pr = fe.pinned_empty(rs.shape); pr[...] = rs
        o1 = hl.alloc_stereo_out(1, pinned=True, full=False, kept=True)
        call = lambda i: hl.check(hl.L.b2orb_stereo_frontend(hl.h, pl[i:].ctypes.data, pr[i:].ctypes.data, 1, w, h, w, w * h,
                                                             1 if temporal else 0, o1[1]))
        lat = []
        for k in range(70):
            t0 = time.perf_counter()
            call(k % npairs)
            lat.append((time.perf_counter() - t0) * 1e3)
        lat = sorted(lat[10:])
        res["latency_p50_ms"] = round(lat[len(lat) // 2], 4)
        res["latency_p95_ms"] = round(lat[int(len(lat) * 0.95)], 4)
        if name == "cfsd":   # config 3: frames released at the IMU rate; the GPU idles between frames (cold clocks / caches)
            period = 1.0 / 200.0
            lat, late = [], 0
            t_start = time.perf_counter() + 0.01
            for k in range(200):
                due = t_start + k * period
                while time.perf_counter() < due:
                    pass
                t0 = time.perf_counter()
                call(k % npairs)
                t1 = time.perf_counter()
                lat.append((t1 - t0) * 1e3)
                late += (t1 - due) > period
            lat = sorted(lat)
            res["paced_200hz"] = {"frames": 200, "period_ms": 5.0, "p50_ms": round(lat[100], 4), "p95_ms": round(lat[190], 4),
                                  "max_ms": round(lat[-1], 4), "deadline_misses": int(late),
                                  "what": "one stereo pair released every 5 ms (busy-wait trigger), latency of the call"}
        hl.close()
        del d_l, d_r, pl, pr
        torch.cuda.empty_cache()
        out[name] = res
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="kitti", help="shape name from synth.SHAPES")
    ap.add_argument("--pairs-per-step", type=int, default=128)
    ap.add_argument("--pool", type=int, default=160, help="distinct synthetic pairs resident (cycled)")
    ap.add_argument("--streams", type=int, default=2, help="handles (CUDA streams) sharing each step on the value leg")
    ap.add_argument("--no-temporal", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--rectify", action="store_true",
                    help="next row f1: inputs are raw camera frames, cv::remap runs on the device in front of the pyramid")
    ap.add_argument("--triangulate", action="store_true",
                    help="next row f4: cv::triangulatePoints + depth gate of the kept matches in the same device pass")
    ap.add_argument("--e2e-threads", type=int, default=3, help="host threads (one handle each) driving the e2e leg")
    ap.add_argument("--bind", action="store_true", help="pin the process to the CPUs local to its GPU also at N=1")
    ap.add_argument("--latency-calls", type=int, default=200, help="single-pair calls of the latency leg")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only: device-resident leg alone (line has e2e null)")
    ap.add_argument("--e2e-out", default="kept", choices=["kept", "full"],
                    help="result set of the e2e leg: what the reference keeps of a frame, or every keypoint/descriptor too")
    ap.add_argument("--no-configs", action="store_true", help="skip the bounded lines for the other BASELINE.json configs")
    ap.add_argument("--debug-set", action="append", default=[], metavar="KEY=VALUE",
                    help="b2orb_debug_set on every handle (A/B runs), e.g. fast_band=2")
    args = ap.parse_args()
    dbg = [(kv.split("=")[0], int(kv.split("=")[1])) for kv in args.debug_set]

    synth = importlib.import_module("pose-estimation_b200.synth")
    w, h, nfeat, ini, mn = synth.SHAPES[args.workload]
    temporal = not args.no_temporal and args.workload.startswith("kitti")
    cfg = (nfeat, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": f"{args.workload} {w}x{h} stereo, {nfeat} ORB feats, 8 levels x1.2, ini/min FAST {ini}/{mn}, "
                          f"BF Hamming stereo{' + temporal circular' if temporal else ''} matching",
              "pairs_per_step": args.pairs_per_step, "streams": args.streams, "pool_pairs": args.pool, "sharding": f"frames x{world} (no collective)",
              "device_input_layout": "u8 rows with a 32-byte aligned pitch (value leg); tight host rows (e2e leg)",
              "l2": "inputs + intermediates per step (>700 MB) exceed the 126 MB L2; the pool of distinct pairs is cycled"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import oracle as orc
        workers = os.cpu_count() or 1
        per_step = args.pairs_per_step          # the same config as the repo arm; a step is one batch of that size
        ls, rs = synth.sequence(args.workload, per_step, seed=1)
        run_pairs, kind = cpu_runner(orc)
        run = lambda: run_pairs(cfg, ls, rs, synth.MATCH_RATIO, synth.MIN_MATCH_DIST, synth.MAX_VERTICAL_PIXEL_DIST,
                                workers, 1, temporal)
        steps, warm = args.steps, args.warmup
        for _ in range(warm):
            run()
        t = time.perf_counter()
        for _ in range(steps):
            run()
        dt = time.perf_counter() - t
        v = per_step * steps / dt
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": workers, "kind": kind,
                             "sample": f"{per_step} pairs/step x {steps} steps, pairs sharded over {workers} host threads",
                             "note": CPU_NOTE},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0

    import torch
    fe = importlib.import_module("pose-estimation_b200.frontend")
    torch.cuda.set_device(local)
    prev_aff, near_cpus = bind_near_gpu(local) if (world > 1 or args.bind) else (None, 0)
    if near_cpus:
        config["host_affinity"] = f"rank threads + pinned buffers on the {near_cpus} CPUs local to the GPU"
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, pool = args.pairs_per_step, max(args.pool, args.pairs_per_step)
    ls, rs = synth.sequence(args.workload, pool, seed=rank)
    pitch = (w + 31) // 32 * 32   # device-resident inputs live in a pitched layout (rows 32-byte aligned), like cudaMallocPitch
    d_l = torch.zeros((pool, h, pitch), dtype=torch.uint8, device="cuda")
    d_r = torch.zeros((pool, h, pitch), dtype=torch.uint8, device="cuda")
    d_l[:, :, :w] = torch.from_numpy(ls).cuda()
    d_r[:, :, :w] = torch.from_numpy(rs).cuda()
    torch.cuda.synchronize()
    # value leg: `nstreams` handles (one CUDA stream each) share the step's pairs, so kernels bound by different
    # resources (ALU-bound FAST, XU-bound Hamming, latency-bound quadtree) overlap on the GPU; per-kernel timing for
    # the roofline comes from a separate single-stream pass afterwards.
    nstr = max(1, min(args.streams, B))
    Bs = B // nstr
    B = Bs * nstr
    def mk_handle(max_images):
        x = fe.Handle(*cfg, w, h, max_images=max_images, device=local)
        for k_, v_ in dbg:
            x.debug_set(k_, v_)
        return x

    hds = [mk_handle(2 * Bs) for _ in range(nstr)]
    streams = [torch.cuda.ExternalStream(hh.stream, device=torch.device("cuda", local)) for hh in hds]
    hd = hds[0]
    rect_maps = None
    if args.rectify:   # + 2 x cv::remap per pair (src/feature-tracker.cpp:75-76), fused into the ingest
        rect_maps = [synth.rectification_maps(w, h, e) for e in (0, 1)]
        for x in hds:
            x.set_rectification(rect_maps[0][0], rect_maps[0][1], rect_maps[1][0], rect_maps[1][1])
        config["rectify"] = "cv::remap INTER_LINEAR, CV_16SC2 maps, fused into the level-0 staging pass"
    proj = None
    if args.triangulate:   # triangulatePoints of the kept matches (src/feature-tracker.cpp:365-379)
        fx = 0.58 * w
        P1 = [[fx, 0, 0.49 * w, 0], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
        P2 = [[fx, 0, 0.49 * w, 0.537 * fx], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]   # synthetic right eye lies towards +x
        proj = (P1, P2, 20.0)
        for x in hds:
            x.set_projection(*proj)
        config["triangulate"] = "cv::triangulatePoints (4x4 Jacobi SVD per kept match, fp64) + depth gates, fused behind the stereo filter"
    img_bytes = w * h

    def step_device(k, only=None):
        off = (k * 37) % (pool - B + 1)
        for i, hh in enumerate(hds if only is None else [only]):
            o = (off + i * Bs) * pitch * h
            hh.stereo_frontend_device(d_l.data_ptr() + o, d_r.data_ptr() + o, Bs, w, h, pitch, pitch * h, temporal)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- value: inputs resident in HBM ----------------
    for k in range(args.warmup):
        step_device(k)
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = sum(hh.launch_count for hh in hds)
    e0 = [torch.cuda.Event(enable_timing=True) for _ in hds]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in hds]
    for i, st in enumerate(streams):
        e0[i].record(st)
    for k in range(args.steps):
        step_device(args.warmup + k)
    for i, st in enumerate(streams):
        e1[i].record(st)
    barrier()
    ms = max(a.elapsed_time(b_) for a in e0 for b_ in e1)   # first start -> last end, on the device
    launches = sum(hh.launch_count for hh in hds) - l0
    tms = torch.tensor([ms], device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    value = world * B * args.steps / (ms_max * 1e-3)
    # per-kernel device times (CUDA events between launches on ONE stream, nothing else running)
    psteps = max(2, min(args.steps, 5))
    hd.profile_enable(True)
    for k in range(psteps):
        step_device(k, only=hd)
    prof = hd.profile_read()
    hd.profile_enable(False)
    prof_pairs = Bs

    # ---------------- e2e: host buffers through the C ABI, copies inside ----------------
    nh = max(args.e2e_threads, 1)
    Be = B                                            # pairs per C-ABI call (one value-leg step), per host thread
    handles = [mk_handle(2 * Be) for _ in range(nh)]
    if rect_maps is not None:
        for x in handles:
            x.set_rectification(rect_maps[0][0], rect_maps[0][1], rect_maps[1][0], rect_maps[1][1])
    if proj is not None:
        for x in handles:
            x.set_projection(*proj)
    pin_l = fe.pinned_empty(ls.shape); pin_l[...] = ls
    pin_r = fe.pinned_empty(rs.shape); pin_r[...] = rs
    cap = handles[0].keypoint_capacity
    e2e_steps = max(args.steps // 2, 4)
    counts = []
    last_off = {}
    errors = []

    def d2h_bytes(mode, n_l, n_r, n_match, n_circ, npairs):
        """Bytes the device writes to host memory for one call, from the counts (what actually crosses the bus)."""
        if mode == "full":   # whole cap-sized slabs (copy path)
            return npairs * (2 * cap * (28 + 32) + 6 * cap * 4 + (cap * (32 + 24 + 1) if proj is not None else 0)) + 16 * npairs
        kept = n_match * (2 * 16 + 2 * 32 + 1 + 2 * 4) + n_circ * 8 + 16 * npairs    # valid prefixes only (export kernel)
        return kept + (n_match * (32 + 24 + 1) if proj is not None else 0)

    def worker(mode, outs_, i, nsteps, record):
        hh, out = handles[i], outs_[i]
        for k in range(nsteps):
            off = ((2 * k + i) * 37) % (pool - Be + 1)
            rc = hh.L.b2orb_stereo_frontend(hh.h, pin_l[off:].ctypes.data, pin_r[off:].ctypes.data, Be, w, h, w, img_bytes,
                                            1 if temporal else 0, out[1])
            if rc != 0:
                errors.append((i, k, rc, hh.L.b2orb_last_error(hh.h).decode()))
                return
            last_off[(mode, i)] = off
            if record:
                o_ = out[0]
                counts.append((int(o_["n_l"].sum()), int(o_["n_match"].sum()), int(o_["n_circ"].sum()), int(o_["n_r"].sum())))

    def run_e2e(mode, outs_, nsteps, record):
        ts = [threading.Thread(target=worker, args=(mode, outs_, i, nsteps, record)) for i in range(nh)]
        for t in ts: t.start()
        for t in ts: t.join()
        if errors:
            raise RuntimeError(f"b2orb_stereo_frontend failed in the e2e leg: {errors[0]}")

    def time_e2e(mode):
        outs_ = [hh.alloc_stereo_out(Be, pinned=True, full=(mode == "full"), kept=True) for hh in handles]
        del counts[:]
        run_e2e(mode, outs_, 2, False)
        barrier()
        t0_ = time.perf_counter()
        run_e2e(mode, outs_, e2e_steps, True)
        torch.cuda.synchronize()
        dt_ = max(time.perf_counter() - t0_, 1e-9)
        te_ = torch.tensor([dt_], device="cuda")
        if world > 1:
            dist.all_reduce(te_, op=dist.ReduceOp.MAX)
        c = counts[-1] if counts else (0, 0, 0, 0)
        return world * nh * Be * e2e_steps / float(te_.item()), d2h_bytes(mode, c[0], c[3], c[1], c[2], Be), outs_, list(counts)

    e2e_value = e2e_d2h = e2e_alt = None
    e2e_outs, e2e_counts = None, []
    if not args.skip_e2e:
        e2e_value, e2e_d2h, e2e_outs, e2e_counts = time_e2e(args.e2e_out)
        if world == 1 and not args.no_configs:   # the other result set, for the comparison (not the headline)
            alt = "full" if args.e2e_out == "kept" else "kept"
            v_alt, b_alt, _, _ = time_e2e(alt)
            e2e_alt = {"out": alt, "value": v_alt, "unit": UNIT, "d2h_bytes_per_step": b_alt}
    clocks = sampler.stop() if sampler else None
    counts = e2e_counts

    # ---------------- p50 frame latency: ONE stereo pair per C-ABI call, pinned host images in, host results out --------
    # (the second half of BASELINE.json's metric; what visual-inertial-slam.cpp sees per frame)
    latency = None
    if not args.skip_e2e:
        hl = mk_handle(2)
        if rect_maps is not None:
            hl.set_rectification(rect_maps[0][0], rect_maps[0][1], rect_maps[1][0], rect_maps[1][1])
        if proj is not None:
            hl.set_projection(*proj)
        out1 = hl.alloc_stereo_out(1, pinned=True)
        lat = []
        for k in range(args.latency_calls + 20):
            i = k % pool
            t0 = time.perf_counter()
            hl.check(hl.L.b2orb_stereo_frontend(hl.h, pin_l[i:].ctypes.data, pin_r[i:].ctypes.data, 1, w, h, img_bytes // h, img_bytes,
                                                1 if temporal else 0, out1[1]))
            lat.append((time.perf_counter() - t0) * 1e3)
        lat = sorted(lat[20:])
        latency = {"p50_ms": round(lat[len(lat) // 2], 4), "p95_ms": round(lat[int(len(lat) * 0.95)], 4),
                   "max_ms": round(lat[-1], 4), "calls": len(lat), "device_ms_last_call": round(hl.last_device_ms, 4),
                   "what": "b2orb_stereo_frontend, 1 pair per call, pinned host images in -> keypoints, descriptors, "
                           "matches on the host; wall clock around the call"}
        hl.close()
        # The UNCHANGED-caller path: what the reference's own call sequence costs with nothing fused —
        # two extractor instances (the two std::threads of src/feature-tracker.cpp:153-156 = two handles, async), PAGEABLE
        # images, then BFMatcher.match on host descriptors, then the host filter loops as b2orb_stereo_filter.
        if rank == 0 and rect_maps is None:
            hL, hR = mk_handle(2), mk_handle(2)
            lat_u = []
            for k in range(min(args.latency_calls, 100) + 10):
                i = k % pool
                il, ir = ls[i], rs[i]                               # pageable numpy arrays, tight rows
                t0 = time.perf_counter()
                hL.extract_async(il); hR.extract_async(ir)
                kl_, dl_ = hL.extract_wait(); kr_, dr_ = hR.extract_wait()
                ti_, di_ = hL.match(dl_, dr_)
                hL.stereo_filter(ti_, di_, kl_, dl_, kr_, dr_)
                lat_u.append((time.perf_counter() - t0) * 1e3)
            lat_u = sorted(lat_u[10:])
            latency["unchanged_caller"] = {
                "p50_ms": round(lat_u[len(lat_u) // 2], 4), "p95_ms": round(lat_u[int(len(lat_u) * 0.95)], 4),
                "what": "the reference's own call sequence, nothing fused: b2orb_extract_async on two handles (pageable images) + "
                        "_wait, b2orb_match (host descriptors in, host lists out), b2orb_stereo_filter; through the ctypes binding, "
                        "which allocates and copies its result arrays per call"}
            hL.close(); hR.close()

    # ---------------- parity spot checks of what was just timed (the oracle is the checker, never the thing measured) --------
    verified = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import oracle as orc_chk
        ex_chk = orc_chk.OracleExtractor(*cfg)

        def check_pair(o_, p_, il_, ir_):
            kl, dl = ex_chk(il_); kr, dr = ex_chk(ir_)
            ti, di = orc_chk.bf_match(dl, dr)
            ml, mr = orc_chk.stereo_filter(ti, di, kl, kr, synth.MATCH_RATIO, synth.MIN_MATCH_DIST, synth.MAX_VERTICAL_PIXEL_DIST)
            m_ = int(o_["n_match"][p_])
            ok = m_ == len(ml) and np.array_equal(o_["match_l"][p_, :m_], ml) and np.array_equal(o_["match_r"][p_, :m_], mr)
            if ok and "kp_l" in o_:
                nl_, nr_ = int(o_["n_l"][p_]), int(o_["n_r"][p_])
                ok = nl_ == len(kl) and nr_ == len(kr) and o_["kp_l"][p_, :nl_].tobytes() == kl.tobytes() and \
                    o_["kp_r"][p_, :nr_].tobytes() == kr.tobytes() and np.array_equal(o_["desc_l"][p_, :nl_], dl) and \
                    np.array_equal(o_["desc_r"][p_, :nr_], dr)
            if ok and "cur_desc_l" in o_:
                ok = np.array_equal(o_["cur_desc_l"][p_, :m_], dl[ml]) and np.array_equal(o_["cur_desc_r"][p_, :m_], dr[mr]) and \
                    np.array_equal(o_["pix_l"][p_, :m_], np.stack([kl["x"][ml], kl["y"][ml]], 1).astype(np.float64))
            return bool(ok)

        verified = {"checker": "CPU oracle, bit-exact compare", "value_leg_pairs": 0, "e2e_leg_pairs": 0, "ok": True}
        if rect_maps is None:
            dl_out = hd.alloc_stereo_out(Bs)
            hd.stereo_download(Bs, dl_out)
            off_v = ((psteps - 1) * 37) % (pool - B + 1)        # the profile pass ran last on this handle
            for p_ in (0, Bs // 2, Bs - 1):
                verified["ok"] &= check_pair(dl_out[0], p_, ls[off_v + p_], rs[off_v + p_])
                verified["value_leg_pairs"] += 1
            if e2e_outs is not None:
                for i_ in range(nh):
                    off_e = last_off.get((args.e2e_out, i_))
                    if off_e is None:
                        continue
                    for p_ in (0, Be - 1):
                        verified["ok"] &= check_pair(e2e_outs[i_][0], p_, ls[off_e + p_], rs[off_e + p_])
                        verified["e2e_leg_pairs"] += 1
        if not verified["ok"]:
            raise RuntimeError("bench: a timed leg's output differs from the oracle")

    # ---------------- the drop-in sequence: one frame per call against the device-resident feature pool (f3) --------------
    # processImage + featurePoolUpdate per frame the way visual-inertial-slam.cpp drives them (keyframe every 3rd frame,
    # maxFeatureAge 8): b2orb_stereo_frontend(temporal = 2) with the small result set, then b2orb_pool_update.
    sequence = None
    if rank == 0 and world == 1 and not args.skip_e2e and rect_maps is None:
        hs = mk_handle(2)
        if proj is not None:
            hs.set_projection(*proj)
        hs.pool_reset(8)
        out_s = hs.alloc_stereo_out(1, pinned=True, full=False, kept=True)
        nfr = min(args.latency_calls, pool)
        lat_s = []
        t_all = time.perf_counter()
        for t_ in range(nfr):
            t0 = time.perf_counter()
            hs.check(hs.L.b2orb_stereo_frontend(hs.h, pin_l[t_:].ctypes.data, pin_r[t_:].ctypes.data, 1, w, h, w, img_bytes, 2, out_s[1]))
            hs.pool_update(t_ % 3 == 0)
            lat_s.append((time.perf_counter() - t0) * 1e3)
        t_all = time.perf_counter() - t_all
        lat_s = sorted(lat_s[5:])
        sequence = {"frames": nfr, "frames_per_s": round(nfr / t_all, 1), "p50_ms": round(lat_s[len(lat_s) // 2], 4),
                    "p95_ms": round(lat_s[int(len(lat_s) * 0.95)], 4), "pool_features_at_end": hs.pool_size,
                    "what": "per frame: b2orb_stereo_frontend(1 pair, temporal=2: externalTrack against the device-resident "
                            "feature pool) + b2orb_pool_update (keyframe every 3rd frame, maxFeatureAge 8); pinned images in, "
                            "kept pixels / descriptors / circular lists / mask out"}
        hs.close()

    # per-GPU result summaries are the only thing gathered (north star: no collective on the hot path)
    summ = torch.tensor([sum(c[0] for c in counts), sum(c[1] for c in counts), sum(c[2] for c in counts)], device="cuda")
    if world > 1:
        gathered = [torch.zeros_like(summ) for _ in range(world)]
        dist.all_gather(gathered, summ)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---------------- roofline of the dominant kernel ----------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    b_pair, per_kernel = algorithmic_bytes(level_sizes(w, h), nfeat, temporal)
    tot_ms = sum(v[0] for v in prof.values())
    dom = max(prof, key=lambda k: prof[k][0])
    shares = {k: round(v[0] / tot_ms, 4) for k, v in prof.items() if v[1]}
    dom_ms_per_launch = prof[dom][0] / max(prof[dom][1], 1)
    launches_per_step = prof[dom][1] / psteps
    # per_kernel is bytes per PAIR (both eyes); one profiled step covers prof_pairs pairs in `launches_per_step` launches
    dom_bytes = per_kernel.get(dom, 0) * prof_pairs / max(launches_per_step, 1)
    achieved = dom_bytes / (dom_ms_per_launch * 1e-3) / 1e9 if dom_ms_per_launch > 0 else 0.0
    # DRAM traffic of that kernel from the committed ncu --set full capture (tools/ncu_traffic.py), scaled from the
    # captured launch to this launch by image count; only valid for the workload it was captured on
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r2", "traffic.json")
    if not os.path.exists(tpath):
        tpath = os.path.join(ROOT, "profiles", "r1", "traffic.json")
    if args.workload == "kitti" and os.path.exists(tpath):
        tj = json.load(open(tpath))
        if dom in tj["kernels"]:
            traffic = tj["kernels"][dom]["dram_bytes_per_image"] * 2 * prof_pairs / max(launches_per_step, 1)
            traffic_src = f"{os.path.relpath(tpath, ROOT)} ({tj['source']}, {tj['images_per_launch']} images/launch, scaled by image count)"
    # instruction-issue roofline: warp instructions per pair (ncu smsp__inst_executed.sum per kernel, committed by
    # tools/ncu_traffic.py) x pairs/s against 4 warp instructions / clk / SM x 148 SMs at the SM clock sampled during the run
    issue = None
    if args.workload == "kitti" and os.path.exists(tpath):
        tj = json.load(open(tpath))
        per_img = {k: v.get("inst_per_image") for k, v in tj["kernels"].items() if v.get("inst_per_image")}
        per_pair_only = {k: v.get("inst_per_pair") for k, v in tj["kernels"].items() if v.get("inst_per_pair")}
        if per_img:
            inst_pair = 2 * sum(per_img.values()) + sum(per_pair_only.values())
            clk = ((clocks or {}).get("sm_mhz") or 1965) * 1e6
            peak_inst = 4 * 148 * clk
            issue = {"inst_per_pair": inst_pair, "peak_inst_per_s": peak_inst, "frac": inst_pair * value / world / peak_inst,
                     "unit": "warp instructions", "source": os.path.relpath(tpath, ROOT),
                     "dominant_kernel_inst_per_pair": 2 * per_img.get(dom, 0) or per_pair_only.get(dom),
                     "note": "mixed ALU+FMA issue measured at 2.9 warp-instr/clk/SM on this part (profiles/r2/pipes.txt): "
                             "0.72 of the 4/clk scheduler limit is the practical ceiling of min/max-dominated code"}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": config,
        "e2e": None if e2e_value is None else {
            "value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * img_bytes * Be, "d2h_bytes_per_step": e2e_d2h,
            "pairs_per_call": Be, "host_threads": nh, "steps": e2e_steps * nh, "out": args.e2e_out,
            "what": "b2orb_stereo_frontend on pinned host images; H2D, kernels and the device->host write of the results inside "
                    "the timed region; 'kept' = _curPixelsL/R, _curDescriptorsL/R, match and circular lists, mask "
                    "(src/feature-tracker.cpp:210-218, 311-321), 'full' adds every keypoint and descriptor of both eyes"},
        "e2e_other_result_set": e2e_alt,
        "latency": latency,
        "sequence": sequence,
        "verified": verified,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "issue": issue, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": dom_ms_per_launch,
                     "note": "the path is byte/integer stencil work that streams each pixel once: HBM is the roofline the "
                             "metric asks for, but every kernel is instruction-issue bound on B200 (see roofline.issue and "
                             "profiles/r2/*_ncu_full_summary.csv, DESIGN.md 4)",
                     "kernel_time_shares": shares,
                     "kernel_us_per_pair": {k: round(v[0] * 1e3 / (psteps * prof_pairs), 3) for k, v in prof.items() if v[1]}},
        "pipeline": {"algorithmic_bytes_per_pair": b_pair, "achieved_gbs": b_pair * value / world / 1e9,
                     "frac_of_hbm_peak": b_pair * value / world / 1e9 / peak},
        "results": {"keypoints_left": int(summ[0]), "stereo_matches": int(summ[1]), "circular_matches": int(summ[2])},
    }
    if prev_aff is not None:
        os.sched_setaffinity(0, prev_aff)
    if not args.no_cpu_baseline and world >= 1:
        from oracle import oracle as orc
        workers = os.cpu_count() or 1
        v, n, dt, kind, stages = cpu_baseline(synth, orc, args.workload, cfg, temporal, workers, args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": workers, "kind": kind,
                                "sample": f"{n} pairs of the same workload in {dt:.1f} s, pairs sharded over {workers} host threads",
                                "stage_ms_per_pair_1_thread": stages, "cv2_primitives_ms": cv2_primitive_ms(synth, args.workload, nfeat),
                                "note": CPU_NOTE}
        cv2p = line["cpu_baseline"]["cv2_primitives_ms"]
        if cv2p:
            # What a SIMD-OpenCV build of the reference would need per pair on one thread, from below: OpenCV's own kernels
            # for the stages OpenCV provides (FAST as whole-level calls: fewer, larger calls than the reference's per-cell
            # ones) + the reference's own code for the rest (quadtree, orientation, descriptors), the cheaper matcher
            ms = cv2p["pyramid"] + cv2p["fast_all_levels_whole_image_calls"] + cv2p["blur"] + stages.get("quadtree", 0.0) + \
                stages.get("orientation", 0.0) + stages.get("describe", 0.0) + \
                min(cv2p["match_stereo"], stages.get("match_stereo", 1e9)) + stages.get("match_temporal", 0.0)
            line["cpu_baseline"]["cv2_hybrid_lower_bound"] = {
                "ms_per_pair_1_thread": round(ms, 2), "pairs_per_s_all_cores_if_linear": round(workers * 1e3 / ms, 1),
                "what": "cv2 4.x SIMD kernels for resize / FAST (whole-level calls) / GaussianBlur + the reference's own "
                        "quadtree, orientation and descriptor code + the faster of cv2.BFMatcher and the restated matcher; "
                        "an estimate from stage times, not a run of the reference"}
        # the reference's own structure (src/feature-tracker.cpp:153-156): pairs one after the other, left and right
        # extraction on two threads — what one instance of the reference does on this host
        fl, fr = synth.sequence(args.workload, 6, seed=1)
        t0 = time.perf_counter()
        cpu_runner(orc)[0](cfg, fl, fr, synth.MATCH_RATIO, synth.MIN_MATCH_DIST, synth.MAX_VERTICAL_PIXEL_DIST, 1, 2, temporal)
        fdt = time.perf_counter() - t0
        if latency is not None:
            latency["cpu_port_ms"] = round(fdt / 6 * 1e3, 2)   # one pair on the host, reference structure (2 threads)
        line["cpu_baseline"]["faithful"] = {"value": 6 / fdt, "unit": UNIT, "cores": 2,
                                            "sample": f"6 pairs sequentially, L/R extraction on 2 threads, {fdt:.2f} s"}
    if world == 1 and not args.no_configs and not args.skip_e2e:
        line["configs"] = other_configs(fe, synth, local, peak, dbg, None if args.no_cpu_baseline else "oracle")
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
