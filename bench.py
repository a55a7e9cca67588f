#!/usr/bin/env python3
"""Open-loop ME benchmark (BASELINE.json metric: "open-loop ME frames/s").

One step = the ME stage (pyramid build, pre-HME, HME L0/L1/L2, full-pel search, candidate
construction) of one mini-GOP of synthetic pictures. Default workload = BASELINE.json configs[1]:
1920x1080 8-bit, preset 4, random access, 240-frame panning-noise clip, 1 B200.

  python bench.py [--gpus N --steps K --warmup W]          our CUDA path (one rank per GPU under torchrun)
  python bench.py --impl reference [...]                   the reference's own CPU ME (oracle/_ref) on the host cores

Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for what each key means."""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

INT_PEAK_LANE_INSTR_PER_CLK_SM = 64.0  # VABSDIFF4.U8.ACC is an ALU-pipe instruction: 64 lane-instr/clk/SM, shared with SHF/LOP3/VIMNMX (profiles/r02_int_peak.md)
N_SM = 148


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--preset", type=int, default=4)
    ap.add_argument("--bits", type=int, default=8, choices=[8, 10])
    ap.add_argument("--frames", type=int, default=240, help="length of the synthetic clip")
    ap.add_argument("--concurrency", type=int, default=2, choices=[1, 2], help="compute streams per session (svtme_session_set_concurrency)")
    ap.add_argument("--e2e-probe", default="full", choices=["full", "h2d", "d2h", "upload"], help="diagnostic: drop one direction of the end-to-end copies; upload = host uploads and pyramid builds only, no ME")
    ap.add_argument("--host-numa", default="auto", choices=["auto", "off"],
                    help="auto: run this rank's threads on (or, failing that, bind its memory to) the NUMA node of its GPU, so that "
                         "the pinned frames and result arrays it allocates are local to the GPU's PCIe root")
    ap.add_argument("--qp", type=int, default=35)
    ap.add_argument("--cpu-baseline-s", type=float, default=12.0, help="wall-time budget of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the short runs of the other BASELINE.json configurations")
    ap.add_argument("--extra-configs", default="4k,8k,m-2,streams", help="which of the other BASELINE.json configurations to run (comma list)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# workload description
# ------------------------------------------------------------------------------------------------
class Workload:
    def __init__(self, a):
        from svt_av1_psy_b200.gop import mrp_for_preset
        self.w, self.h, self.preset, self.bits, self.n_frames, self.qp = a.width, a.height, a.preset, a.bits, a.frames, a.qp
        # enc_handle.c:4791-4799: 5 hierarchical levels up to M9 below 8K, else 4
        self.hier = 4 if (self.w * self.h >= 0x140A000 or self.preset > 9 or (self.preset >= 9 and self.w * self.h >= 0x16DA00)) else 5
        self.mgop = 1 << self.hier
        self.mrp = mrp_for_preset(self.preset)
        self.name = "%dx%d %d-bit preset %d random-access %dL, %d-frame panning-noise clip" % (
            self.w, self.h, self.bits, self.preset, self.hier, self.n_frames)

    def plan(self, poc):
        from svt_av1_psy_b200 import gop
        return gop.picture_plan(poc, self.hier, self.mrp["l0"], self.mrp["l1"], 1 << 30)

    def order(self, first_poc):
        from svt_av1_psy_b200 import gop
        return gop.coding_order(first_poc, self.hier, 1 << 30)


def make_clip(wl, rank):
    """n_frames distinct frames; content of picture `poc` is frame poc % n_frames (the POC structure keeps running)"""
    from svt_av1_psy_b200.synth import PanningNoise
    clip = PanningNoise(wl.w, wl.h, seed=7 + rank, vx=2, vy=1, max_frames=wl.n_frames, sigma=2.0)
    rng = np.random.default_rng(1234 + rank)
    # cheap per-frame noise: crops of one pre-generated noise field (instead of a fresh normal draw per frame)
    nf = np.rint(rng.normal(0.0, clip.sigma, size=(wl.h + 64, wl.w + 64))).astype(np.int16)
    frames = np.empty((wl.n_frames, wl.h, wl.w), np.uint8)
    for t in range(wl.n_frames):
        x0, y0 = abs(clip.vx) * t, abs(clip.vy) * t
        oy, ox = rng.integers(0, 64, 2)
        crop = clip.field[y0:y0 + wl.h, x0:x0 + wl.w].astype(np.int16) + nf[oy:oy + wl.h, ox:ox + wl.w]
        np.clip(crop, 0, 255, out=crop)
        frames[t] = crop.astype(np.uint8)
    return frames


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.windows = index, [], None, []

    def start(self):
        """nvidia-smi takes ~1 s to produce its first line: started once, before the warm-up; samples are
        stamped on arrival and only those that fall inside a timed window are reported."""
        if os.environ.get("SVTME_NO_SAMPLER"):  # diagnostic: is the polling itself disturbing the copies?
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        ok = [(t, r) for t, r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        inside = [r for t, r in ok if any(a <= t <= b for a, b in self.windows)]
        where = "timed regions"
        if len(inside) < 3:  # a very short timed region (small --steps): fall back to every sample taken under load
            t_first = min((a for a, _ in self.windows), default=0) - 2.0
            inside, where = [r for t, r in ok if t >= t_first], "warm-up + timed regions (timed region shorter than 3 samples)"
        sm = sorted(float(r[0]) for r in inside)
        mx = [float(r[1]) for r in inside if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in inside for i in range(4) if r[3 + i].lower().startswith("active")})
        pw = sorted(float(r[2]) for r in inside if r[2].replace(".", "").isdigit())
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "window": where, "power_w": pw[len(pw) // 2] if pw else None}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
class StepPlan:
    """Everything one step hands to the C ABI, marshalled ahead of the timed region: picture ids, parameter array, the
    upload lists (host pointers of this step's new frames). An encoder's host code is C and keeps these arrays; building
    them in Python inside the timed loop would time the harness, not the library."""
    __slots__ = ("n", "ids", "prm", "uploads", "h2d")


class Runner:
    """Drives one Session through consecutive mini-GOPs of a stream (this rank's shard)."""
    N_SLOTS = 192  # picture store: slot = poc % N_SLOTS; references reach at most 4 mini-GOPs back

    def __init__(self, wl, rank, world, frames_host, device, fields, own_stream=False):
        import torch
        import svtme_b200 as me
        from svt_av1_psy_b200 import abi
        self.wl, self.rank, self.world, self.me, self.abi, self.torch = wl, rank, world, me, abi, torch
        self.sess = me.Session(wl.w, wl.h, max_pictures=self.N_SLOTS, device=device)
        if not own_stream:  # own_stream: the session keeps the compute stream it created (several sessions run concurrently on the GPU)
            self.sess.set_stream(torch.cuda.current_stream().cuda_stream)
        self.frames_pinned = torch.from_numpy(frames_host).pin_memory()
        self.frames_dev = self.frames_pinned.to("cuda", non_blocking=False)
        self.frames16 = None
        if wl.bits == 10:  # host side holds 10-bit samples (uint16); ME reads their 8 MSBs = frames_host
            rng = np.random.default_rng(99 + rank)
            f16 = (frames_host.astype(np.uint16) << 2) | rng.integers(0, 4, frames_host.shape, dtype=np.uint16)
            self.frames16 = torch.from_numpy(f16.view(np.int16)).pin_memory()
        self.resident = {}  # slot -> poc whose planes are in the store (as of the last PLANNED step)
        self.params_cache = {}
        self.fields = fields
        self.sess.set_result_fields(fields)  # what the kernels produce when results stay in HBM: the same arrays
        self.slabs = None
        self.h2d = self.d2h = 0
        self.pics = 0
        self.host_batches = 0
        self.call_s = 0.0
        self.upload16 = True  # 10-bit workloads: upload the 16-bit samples (split on the device) / the 8-MSB plane the encoder keeps
        self.probe = "full"

    def params(self, plan):
        key = plan["poc"]
        sin = self.me.binding.SigDerivIn()
        wl = self.wl
        sin.enc_mode, sin.width, sin.height, sin.qp = wl.preset, wl.w, wl.h, wl.qp
        sin.pred_structure, sin.hierarchical_levels = 2, wl.hier
        sin.temporal_layer_index, sin.is_ref, sin.sc_class1 = plan["tl"], plan["is_ref"], 0
        sin.frame_rate, sin.similar_brightness_refs, sin.gm_level = 30 << 16, 0, -1
        sin.only_l_bwd, sin.safe_limit_nref, sin.safe_limit_zz_th = wl.mrp["only_l_bwd"], wl.mrp["safe_limit_nref"], wl.mrp["safe_limit_zz_th"]
        sin.max_l0, sin.max_l1 = wl.mrp["l0"], wl.mrp["l1"]
        sin.picture_number, sin.num_lists = key, 2
        lists = [plan["l0"], plan["l1"]]
        for l in range(2):
            sin.ref_count[l] = len(lists[l])
            for r, p in enumerate(lists[l]):
                sin.ref_picture_number[l][r] = p
        p = self.me.sig_deriv_me(sin)
        diag = os.environ.get("SVTME_DIAG_PREHME")  # diagnostic only (NOT the reference's configuration): shrink one pre-HME region
        if diag:
            for i in range(2):
                if str(i) in diag:
                    a = p.ctrls.prehme_sa_cfg[i]
                    a.sa_min.width, a.sa_min.height, a.sa_max.width, a.sa_max.height = 8, 3, 8, 3
        for l in range(2):
            for r, rp in enumerate(lists[l]):
                p.ref_pic_id[l][r] = rp % self.N_SLOTS
        return p

    def plan_step(self, step_index, host):
        """mini-GOP m = step*world + rank: its pictures (coding order), the frames the picture store does not hold yet"""
        wl = self.wl
        m = step_index * self.world + self.rank
        first = m * wl.mgop
        pending = {}
        sp = StepPlan()
        sp.h2d = 0

        def ensure(poc):
            slot = poc % self.N_SLOTS
            if self.resident.get(slot) == poc:
                return
            f = poc % wl.n_frames
            if host and self.probe == "d2h":
                kind, ptr = "device", self.frames_dev[f].data_ptr()
            elif host and self.frames16 is not None and self.upload16:
                kind, ptr = "host10", self.frames16[f].data_ptr()
                sp.h2d += 2 * wl.w * wl.h
            elif host:
                kind, ptr = "host8", self.frames_pinned[f].data_ptr()
                sp.h2d += wl.w * wl.h
            else:
                kind, ptr = "device", self.frames_dev[f].data_ptr()
            pending.setdefault(kind, ([], []))
            pending[kind][0].append(slot)
            pending[kind][1].append(ptr)
            self.resident[slot] = poc

        ensure(first)  # the base/key picture that precedes the mini-GOP
        ids, prms = [], []
        for poc in wl.order(first):
            plan = wl.plan(poc)
            for p in plan["l0"] + plan["l1"]:
                ensure(p)
            ensure(poc)
            prm = self.params_cache.get(poc)
            if prm is None:
                prm = self.params_cache[poc] = self.params(plan)
            ids.append(poc % self.N_SLOTS)
            prms.append(prm)
        sp.n = len(ids)
        sp.ids = (C.c_int32 * sp.n)(*ids)
        sp.prm = (self.abi.MePicParams * sp.n)(*prms)
        sp.uploads = []
        for kind, (slots, ptrs) in pending.items():
            for i in range(0, len(slots), 32):
                k = len(slots[i:i + 32])
                sp.uploads.append((kind, k, (C.c_int32 * k)(*slots[i:i + 32]), (C.c_void_p * k)(*ptrs[i:i + 32])))
        return sp

    def setup_slabs(self, sp):
        """three pinned result slabs used in turn: the host queues batch k + 2 (which reuses the device region of batch k, the
        library waits for that download on the device, before the batch's first writing kernel) without waiting for batch k's
        download itself; it waits only for the batch that last used the slab it is about to hand out again (batch k - 1). With two slabs the host had
        to wait for the download of batch k before it could queue batch k + 2, and the second compute stream ran dry."""
        import torch
        lay = self.sess.packed_layout(sp.prm[0], self.fields)
        self.slab_bytes = lay.bytes * self.wl.mgop
        self.slabs = [torch.zeros(self.slab_bytes, dtype=torch.uint8).pin_memory() for _ in range(3)]
        self.block_bytes = lay.bytes
        self.seqs = [0, 0, 0]
        self.sess.reserve(sp.prm[0], self.wl.mgop)

    def run_step(self, sp, host):
        """the C calls of one step: batched uploads, ONE launch sequence, (host) ONE download of the whole batch"""
        L = self.me.lib()
        h = self.sess._h
        t_call = time.perf_counter()
        for kind, k, slots, ptrs in sp.uploads:
            fn = L.svtme_picture_upload_batch if kind == "host8" else (L.svtme_picture_upload_10bit_batch if kind == "host10" else L.svtme_picture_upload_device_batch)
            rc = fn(h, k, slots, ptrs, self.wl.w)
            if rc:
                raise RuntimeError("upload failed 0x%x" % rc)
        self.h2d += sp.h2d
        if host and self.probe == "upload":
            self.pics += sp.n
            self.call_s += time.perf_counter() - t_call
            return
        if host and self.probe != "h2d":
            k = self.host_batches % 3
            t_wait = time.perf_counter()
            self.sess.wait_batch(self.seqs[k])  # this slab's previous results are in host memory (and consumed) before reuse
            t_call += time.perf_counter() - t_wait  # waiting for the pipeline is not host work
            self.sess.me_batch_packed(sp.ids, sp.prm, sp.n, self.fields, self.slabs[k].data_ptr(), self.slab_bytes)
            self.seqs[k] = self.sess.batch_seq()
            self.host_batches += 1
            self.d2h += self.block_bytes * sp.n
        else:
            rc = L.svtme_picture_me_batch(h, sp.n, sp.ids, sp.prm, None, 0)
            if rc:
                raise RuntimeError("svtme_picture_me_batch failed 0x%x" % rc)
        self.pics += sp.n
        self.call_s += time.perf_counter() - t_call  # host time inside the library's entry points (queueing the step)

    def step(self, step_index, host):
        sp = self.plan_step(step_index, host)
        if not getattr(self, "reserved", False):  # queues, result regions and job arrays sized once, before anything is timed
            self.sess.reserve(sp.prm[0], self.wl.mgop)
            self.reserved = True
        if host and self.slabs is None:
            self.setup_slabs(sp)
        self.run_step(sp, host)


def _cpulist(text):
    out = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        out.update(range(int(lo), int(hi or lo) + 1))
    return out


def bind_to_gpu_numa_node(torch, dev):
    """Host placement for one-process-per-GPU runs: pinned memory is first-touch, so where this process runs decides which
    socket's memory (and whether the inter-socket link) every H2D / D2H copy of the end-to-end pass crosses. Returns what
    was done, for the JSON line."""
    info = {"gpu_node": None, "policy": "none"}
    try:
        p = torch.cuda.get_device_properties(dev)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        info["gpu_node"] = node
        if node < 0:
            return info
        cpus = _cpulist(open("/sys/devices/system/node/node%d/cpulist" % node).read()) & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["policy"] = "cpu affinity (%d cpus of node %d)" % (len(cpus), node)
        else:  # the cpus this container may use are on another socket: bind the memory instead
            import ctypes
            mask = ctypes.c_ulong(1 << node)
            rc = ctypes.CDLL(None, use_errno=True).syscall(238, 2, ctypes.byref(mask), 64)  # set_mempolicy(MPOL_BIND)
            info["policy"] = "membind node %d" % node if rc == 0 else "none (set_mempolicy errno %d)" % ctypes.get_errno()
    except Exception as e:  # no sysfs / no permission: leave the placement to the OS
        info["policy"] = "none (%s)" % type(e).__name__
    return info


def read_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        return {}


def measure(wl, frames, rank, world, local, a, steps, warmup, sampler, detail):
    """resident pass + per-kernel timing pass + end-to-end pass(es) of one workload -> dict (valid on every rank; the
    times are the max over ranks). detail: also the temporal-filter ME and detector lines (the headline workload)."""
    import torch
    import torch.distributed as dist
    from svt_av1_psy_b200 import abi
    consumer = abi.F_SB_RESULTS | abi.F_STATS  # what open-loop ME's consumers read (include/svtme_b200.h)
    run = Runner(wl, rank, world, frames, local, consumer)
    run.probe = a.e2e_probe

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(host, first_step, n_steps, smp=None):
        plans = [run.plan_step(s, host) for s in range(first_step, first_step + n_steps)]  # marshalled outside the timed region
        if host and run.slabs is None:
            run.setup_slabs(plans[0])
        barrier()
        run.pics = run.h2d = run.d2h = 0
        run.call_s = 0.0
        l0 = run.sess.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for sp in plans:
            run.run_step(sp, host)
        run.host_ms = (time.time() - t0) * 1e3 / n_steps  # host wall time per step of the queueing loop (includes waiting for a free result slab / job array when the pipeline is full)
        run.host_call_ms = run.call_s * 1e3 / n_steps     # of which inside the library's entry points, waits excluded
        run.sess.sync()  # every stream of the session: uploads, kernels (two compute streams) and downloads are done
        e1.record()
        barrier()
        if smp:
            smp.window(t0, time.time())
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, run.pics, run.sess.launch_count() - l0

    # ---- device-resident pass: raw frames already in HBM, results stay in HBM ----
    for s in range(warmup):
        run.step(s, host=False)
    run.sess.set_concurrency(a.concurrency)  # 2: consecutive mini-GOPs on two compute streams
    ms, pics, launches = timed(False, warmup, steps, sampler)
    out = {"value": world * pics / (ms * 1e-3), "ms_per_step": ms / steps, "launches": launches, "host_queue_ms": run.host_ms, "pics": pics, "ms": ms}
    # per-kernel times and the |a-b| operation count for the roofline: a short pass with CUDA events behind every launch
    # (the session then keeps its batches on one stream, so the kernels of different mini-GOPs do not overlap)
    run.sess.enable_timing(True)
    run.sess.kernel_time(reset=True)
    run.sess.kernel_times(reset=True)
    run.sess.sad_ops_by_stage(reset=True)
    t_steps = min(steps, 12)
    _, n_timed, _ = timed(False, warmup + steps, t_steps)
    fp_ms, hme_ms, _ = run.sess.kernel_time(reset=True)
    out["kernel_ms"] = run.sess.kernel_times(reset=True)
    out["ops"] = run.sess.sad_ops_by_stage(reset=True)
    out["n_timed"], out["hme_ms"], out["fp_ms"] = n_timed, hme_ms, fp_ms
    run.sess.enable_timing(False)

    # ---- end-to-end pass: pinned host frames in, the consumers' result arrays out, every step ----
    def e2e_pass(fields, base):
        run.resident.clear()
        run.fields, run.slabs = fields, None
        run.sess.set_result_fields(fields)
        for s in range(base, base + min(warmup, 2)):
            run.step(s, host=True)
        base += min(warmup, 2)
        ms_e, pics_e, _ = timed(True, base, steps, sampler)
        return {"value": world * pics_e / (ms_e * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": int(run.h2d // max(1, steps)),
                "d2h_bytes_per_step": int(run.d2h // max(1, steps)), "ms_per_step": ms_e / steps, "host_queue_ms_per_step": run.host_ms,
                "host_call_ms_per_step": run.host_call_ms}, base + steps

    base = warmup + steps + t_steps
    # the two MeSbResults arrays cross PCIe at the picture's own strides (SVTME_F_SB_COMPACT: 8 instead of 23 candidate bytes
    # per PU at this preset); a consumer expands a b64's rows with svtme_sb_expand_b64 when it copies them into pa_me_data
    out["e2e"], base = e2e_pass(consumer | abi.F_SB_COMPACT, base)
    if wl.bits == 10:
        # the encoder itself hands ME the 8-MSB plane of a 10-bit picture (its PA reference buffers are 8-bit, reference_object.c:327):
        # the same pass with that plane as the host input instead of the 16-bit samples (half the H2D bytes)
        run.upload16 = False
        out["e2e_msb_planes"], base = e2e_pass(consumer | abi.F_SB_COMPACT, base)
        run.upload16 = True
    if detail:  # the same with the MeContext members (p_sb_best_sad / p_sb_best_mv, search_results) crossing PCIe too
        out["e2e_all"], base = e2e_pass(consumer | abi.F_BEST | abi.F_SEARCH_RESULTS, base)
        run.sess.set_result_fields(consumer)
    out["run"] = run
    return out


def run_ours(a):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device. The ME path has no CPU fallback (use --impl reference for the CPU arm).")
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(torch, local) if a.host_numa == "auto" else {"gpu_node": None, "policy": "off"}
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = Workload(a)
    frames = make_clip(wl, rank)
    sampler = ClockSampler(local)
    sampler.start()
    m = measure(wl, frames, rank, world, local, a, a.steps, a.warmup, sampler, detail=True)
    clocks = sampler.stop()
    run = m["run"]
    value, e2e = m["value"], m["e2e"]
    n_timed, hme_ms, fp_ms, ops = m["n_timed"], m["hme_ms"], m["fp_ms"], m["ops"]
    hme_ops, fp_ops = ops[0], ops[1]

    # ---- dynamic-GOP detector HME: pairs of resident pictures, metrics back to the host each call ----
    dg_pairs = 64
    run.sess.sync()
    slots = sorted(run.resident.keys())
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(4):
        run.sess.dg_detector(slots[i % len(slots)], slots[(i + 1) % len(slots)], per_b64=False)
    d0.record()
    for i in range(dg_pairs):
        run.sess.dg_detector(slots[i % len(slots)], slots[(i + 3) % len(slots)], per_b64=False)
    d1.record()
    torch.cuda.synchronize()
    dg_ms = d0.elapsed_time(d1) / dg_pairs

    # ---- temporal-filter ME: (centre, neighbour) pairs of resident pictures, batches of 32 ----
    from svt_av1_psy_b200 import abi, binding as _b
    tf_level, tf_batches = 1, 6
    tf_ids = [slots[i % len(slots)] for i in range(32)]
    tf_prms = []
    for i, sid in enumerate(tf_ids):
        p = _b.sig_deriv_me_tf(wl.preset, wl.w, wl.h, tf_level, qp_opt=1, me_exit_th=0, qp=wl.qp, centre_poc=100 + i, ref_poc=99 + i)
        p.ref_pic_id[0][0] = slots[(i + 1) % len(slots)]
        tf_prms.append(p)
    run.sess.set_result_fields(abi.F_BEST | abi.F_SEARCH_RESULTS)  # what the temporal filter reads (temporal_filtering.c:1860-2226, :2672)
    run.sess.me_batch(tf_ids, tf_prms, None, sync=True)
    run.sess.sad_ops(reset=True)
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0e.record()
    for _ in range(tf_batches):
        run.sess.me_batch(tf_ids, tf_prms, None, sync=False)
    run.sess.sync()
    t1e.record()
    torch.cuda.synchronize()
    tf_ms = t0e.elapsed_time(t1e) / (tf_batches * 32)
    tf_ops = sum(run.sess.sad_ops(reset=True)) / (tf_batches * 32)
    del run, m["run"]

    # ---- the other BASELINE.json configurations (each a short run of its own; see extra_configs) ----
    extra = extra_configs(a, rank, world, local) if not a.no_extra_configs else {}

    if rank == 0:
        peaks = read_peaks()
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        n_refs = wl.mrp["l0"] + wl.mrp["l1"]
        n_b64 = ((wl.w + 63) // 64) * ((wl.h + 63) // 64)
        sm_mhz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        int_peak = INT_PEAK_LANE_INSTR_PER_CLK_SM * 4 * N_SM * sm_mhz * 1e6  # |a-b| ops/s at the sampled clock
        # the dominant KERNEL of the sequence by its own CUDA-event time, and the |a-b| operations it executed
        kms = m["kernel_ms"]
        kernel_ops = {"k_search16 (pre-HME)": ops[2], "k_search16 (HME level 0)": ops[3], "k_search_blk (HME level 1)": ops[4],
                      "k_search_blk (HME level 2)": ops[5], "k_fp": fp_ops, "k_fp_big": 0}
        dom = max(kms, key=lambda k: kms[k]) if kms else None
        dom_ms = kms.get(dom, 0.0)
        dom_ops = kernel_ops.get(dom, 0)
        if dom in ("k_fp", "k_fp_big"):  # one search split over two kernels: their times together
            dom, dom_ms = "k_fp + k_fp_big", kms.get("k_fp", 0) + kms.get("k_fp_big", 0)
        n_launch = max(1, n_timed // wl.mgop)  # launches of the dominant kernel in the timed pass = mini-GOPs
        # algorithmic bytes of one launch of that kernel (one mini-GOP): the planes it reads, once each, + what it writes
        lvl = {"k_search16 (pre-HME)": 1 / 16, "k_search16 (HME level 0)": 1 / 16, "k_search_blk (HME level 1)": 1 / 4}.get(dom, 1.0)
        alg_bytes = wl.mgop * ((1 + n_refs) * lvl * wl.w * wl.h + n_b64 * 8 * 4 * 40)  # + the job records (keys) of a launch
        avg_s = dom_ms * 1e-3 / max(1, n_launch)
        traffic = ncu_traffic(dom, wl)
        line = {
            "metric": "open-loop ME frames/s", "value": round(value, 2), "unit": "frames/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(m["ms_per_step"], 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl.name, "pictures_per_step": wl.mgop, "batch": "one launch sequence per mini-GOP (svtme_picture_me_batch), consecutive mini-GOPs on %d compute stream(s); end to end: three pinned result slabs in rotation" % a.concurrency, "sharding": "mini-GOP m -> rank m % N",
                       "results": "the arrays open-loop ME's consumers read: MeSbResults (total_me_candidate_index, me_mv_array, me_candidate_array) + the eight per-b64 PCS statistics; "
                                  "end to end their rows cross PCIe at the picture's own strides (SVTME_F_SB_COMPACT, expanded per b64 by svtme_sb_expand_b64); "
                                  "e2e_all_fields is the same pass at the reference's strides plus the MeContext members (p_sb_best_sad / p_sb_best_mv, search_results), which only the temporal filter reads after the call",
                       "e2e_pipeline": "host uploads, kernels and result downloads on three streams; three host result slabs in rotation (up to three batches queued); per step the new frames cross as merged H2D copies and the whole batch's results land in pinned host memory by ONE D2H copy (svtme_picture_me_batch_packed), inside the timed region",
                       "host_numa": numa,
                       "l2": "inputs larger than L2: %d distinct %dx%d frames (%.0f MB) stream through the picture store, no explicit flush"
                             % (wl.n_frames, wl.w, wl.h, wl.n_frames * wl.w * wl.h / 1e6)},
            "e2e": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in e2e.items()},
            "e2e_all_fields": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in m["e2e_all"].items()},
            "host_queue_ms_per_step": round(m["host_queue_ms"], 3),
            "gpu_launches": int(m["launches"]),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": round(alg_bytes / avg_s / 1e9, 2) if avg_s else None, "peak": hbm_peak,
                         "unit": "GB/s", "frac": round(alg_bytes / avg_s / 1e9 / hbm_peak, 5) if avg_s else None, "traffic": traffic,
                         "avg_launch_ms": round(avg_s * 1e3, 4), "launch": "one mini-GOP of %d pictures" % wl.mgop,
                         "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                         "note": "integer-ALU bound kernel (byte SAD); see roofline_int"},
            "roofline_int": {"bound": "int_alu(VABSDIFF4.U8.ACC)", "kernel": dom,
                             "achieved": round(dom_ops / (dom_ms * 1e-3) / 1e12, 3) if dom_ms else None,
                             "peak": round(int_peak / 1e12, 2), "unit": "T |a-b| ops/s",
                             "frac": round(dom_ops / (dom_ms * 1e-3) / int_peak, 4) if dom_ms else None,
                             "peak_def": "64 lane-instr/clk/SM on the ALU pipe (profiles/r02_int_peak.md: VABSDIFF4, SHF, LOP3, VIMNMX share it; ncu pipe counts) x 4 px x 148 SM x sampled SM clock",
                             "stage_frac": {"hme_stage": round(hme_ops / (hme_ms * 1e-3) / int_peak, 4) if hme_ms else None,
                                            "fullpel_stage": round(fp_ops / (fp_ms * 1e-3) / int_peak, 4) if fp_ms else None,
                                            "whole_step": round((hme_ops + fp_ops) / max(1, n_timed) * m["pics"] / (m["ms"] * 1e-3) / int_peak, 4)}},
            "kernel_ms_per_picture": {k: round(v / max(1, n_timed), 5) for k, v in kms.items()},
            "kernel_frac_of_int_peak": dict(
                [(k, round(kernel_ops[k] / (kms[k] * 1e-3) / int_peak, 4)) for k in kms if k not in ("k_fp", "k_fp_big") and kernel_ops.get(k) and kms[k]] +
                [("k_fp + k_fp_big", round(fp_ops / ((kms.get("k_fp", 0) + kms.get("k_fp_big", 0)) * 1e-3) / int_peak, 4))] if kms.get("k_fp") else []),
            "stage_ms_per_picture": {"hme_stage": round(hme_ms / max(1, n_timed), 4), "fullpel_stage": round(fp_ms / max(1, n_timed), 4)},
            "sad_gops": round((hme_ops + fp_ops) / max(1, n_timed) * m["pics"] / (m["ms"] * 1e-3) / 1e9, 1),
            "tf_me": {"metric": "temporal-filter ME (ME_MCTF, tf hme_me_level %d, FULL_SAD) picture pairs/s, resident" % tf_level,
                      "value": round(1e3 / tf_ms, 1), "ms_per_pair": round(tf_ms, 4), "sad_gops": round(tf_ops / (tf_ms * 1e-3) / 1e9, 1)},
            "dg_detector": {"metric": "dynamic-GOP detector HME picture pairs/s (svtme_dg_detector_hme, synchronous, metrics to host)",
                            "value": round(1e3 / dg_ms, 1), "ms_per_pair": round(dg_ms, 4),
                            "sad_gops_nominal_window": round(n_b64 * 128 * 128 * 256 / (dg_ms * 1e-3) / 1e9, 1) if wl.w * wl.h >= 0xA1400 else None},
            "configs": extra,
        }
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(wl, frames, a.cpu_baseline_s)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ncu_traffic(kernel, wl):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture of this
    workload (profiles/r02_ncu_full_1080p_m4.json, written by tools/ncu_summary.py from the .ncu-rep); None when there is
    no capture for the workload or the kernel."""
    if (wl.w, wl.h, wl.preset, wl.bits) != (1920, 1080, 4, 8) or not kernel:
        return None
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_full_1080p_m4.json")))
    except (OSError, ValueError):
        return None
    names = [kernel] if kernel != "k_fp + k_fp_big" else ["k_fp", "k_fp_big"]
    tot = 0
    for n in names:
        if n not in cap.get("kernels", {}):
            return None
        tot += cap["kernels"][n]["dram_bytes"]
    return int(tot)


def extra_configs(a, rank, world, local):
    """BASELINE.json configs[2..4] as short runs inside the same bench invocation (the driver only ever runs bench.py
    --gpus N): 3840x2160 10-bit preset 4 at every N; at N = 1 also 7680x4320 preset 2 and 1920x1080 preset -2; and the
    16 concurrent 1080p streams of configs[4], 16 / N per rank, one session each. Shorter clips than the headline run
    (still larger than L2); per-config values are whole-job aggregates like the headline."""
    import copy
    import torch
    out = {}
    peaks = read_peaks()

    def one(name, w, h, bits, preset, frames_n, steps, warmup):
        b = copy.copy(a)
        b.width, b.height, b.bits, b.preset, b.frames = w, h, bits, preset, frames_n
        wl = Workload(b)
        frames = make_clip(wl, rank)
        smp = ClockSampler(local)
        smp.start()
        m = measure(wl, frames, rank, world, local, b, steps, warmup, smp, detail=False)
        clocks = smp.stop()
        del m["run"]
        torch.cuda.empty_cache()
        sm_mhz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        int_peak = INT_PEAK_LANE_INSTR_PER_CLK_SM * 4 * N_SM * sm_mhz * 1e6
        ops, n_timed = m["ops"], max(1, m["n_timed"])
        out[name] = {
            "workload": wl.name, "value": round(m["value"], 2), "unit": "frames/s", "steps": steps, "warmup": warmup,
            "pictures_per_step": wl.mgop, "ms_per_step": round(m["ms_per_step"], 4),
            "e2e": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in m["e2e"].items()},
            **({"e2e_msb_planes": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in m["e2e_msb_planes"].items()}} if "e2e_msb_planes" in m else {}),
            "roofline_int": {"whole_step_frac": round((ops[0] + ops[1]) / n_timed * m["pics"] / (m["ms"] * 1e-3) / int_peak, 4),
                             "hme_stage_frac": round(ops[0] / (m["hme_ms"] * 1e-3) / int_peak, 4) if m["hme_ms"] else None,
                             "fullpel_stage_frac": round(ops[1] / (m["fp_ms"] * 1e-3) / int_peak, 4) if m["fp_ms"] else None,
                             "peak": round(int_peak / 1e12, 2), "unit": "T |a-b| ops/s"},
            "sad_gops": round((ops[0] + ops[1]) / n_timed * m["pics"] / (m["ms"] * 1e-3) / 1e9, 1),
            "stage_ms_per_picture": {"hme_stage": round(m["hme_ms"] / n_timed, 4), "fullpel_stage": round(m["fp_ms"] / n_timed, 4)},
            "kernel_ms_per_picture": {k: round(v / n_timed, 5) for k, v in m["kernel_ms"].items()},
            "clocks": clocks,
        }

    want = set(a.extra_configs.split(","))
    if 16 % world == 0 and "streams" in want:  # first: sixteen sessions are created before the large configurations have churned device memory
        out["16 concurrent 1920x1080 streams"] = streams16(a, rank, world, local)
    if "4k" in want:
        one("3840x2160 10-bit preset 4", 3840, 2160, 10, 4, 96, 6, 3)
    if world == 1 and "8k" in want:
        one("7680x4320 8-bit preset 2", 7680, 4320, 8, 2, 48, 3, 3)
    if world == 1 and "m-2" in want:
        one("1920x1080 8-bit preset -2", 1920, 1080, 8, -2, 96, 3, 3)
    return out


def streams16(a, rank, world, local):
    """configs[4]: 16 concurrent 1080p preset-4 streams over N GPUs, 16 / N sessions per rank, mini-GOPs of the streams
    submitted round-robin (each session has its own three CUDA streams: the streams overlap on the GPU)."""
    import copy
    import torch
    import torch.distributed as dist
    from svt_av1_psy_b200 import abi
    per_rank = 16 // world
    b = copy.copy(a)
    b.width, b.height, b.bits, b.preset, b.frames = 1920, 1080, 8, 4, 64
    wl = Workload(b)
    fields = abi.F_SB_RESULTS | abi.F_SB_COMPACT | abi.F_STATS
    runs = [Runner(wl, 0, 1, make_clip(wl, 100 + rank * per_rank + k), local, fields, own_stream=True) for k in range(per_rank)]
    steps, warmup = 3, 3
    res = {}
    smp = ClockSampler(local)
    smp.start()
    for host in (False, True):
        for r in runs:
            r.resident.clear()
        for s in range(warmup):
            for r in runs:
                r.step(s + (100 if host else 0), host)
        plans = [[r.plan_step(warmup + s + (100 if host else 0), host) for r in runs] for s in range(steps)]
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        for r in runs:
            r.pics = r.h2d = r.d2h = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for ps in plans:
            for r, sp in zip(runs, ps):
                r.run_step(sp, host)
        for r in runs:
            r.sess.sync()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        smp.window(t0, time.time())
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        pics = sum(r.pics for r in runs) * world
        res["e2e" if host else "resident"] = {"value": round(pics / (ms * 1e-3), 2), "unit": "frames/s (all 16 streams)", "ms": round(ms, 3)}
        if host:
            res["e2e"]["h2d_bytes_per_step"] = int(sum(r.h2d for r in runs) // steps)
            res["e2e"]["d2h_bytes_per_step"] = int(sum(r.d2h for r in runs) // steps)
    res["clocks"] = smp.stop()
    res["workload"] = "16 streams of %s, %d sessions per rank x %d ranks, %d mini-GOPs per stream timed" % (wl.name, per_rank, world, steps)
    for r in runs:
        r.sess.close()
    torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own ME on the host cores (oracle/_ref), else the oracle port
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """The reference's ME stage on the host cores, one picture per thread, with the table its x86 build installs
    (aom_dsp_rtcd.c:499,507-514: AVX-512 SAD loop + AVX2 block kernels; AVX2 only without avx512bw). The harness's plane
    cache is on, so a picture's pyramid is built once (inside the timed region, as in our arm) and not once per use as a
    reference. Test infrastructure used as the timed CPU baseline only."""

    def __init__(self, wl, frames, threads):
        from concurrent.futures import ThreadPoolExecutor
        from oracle import pyoracle as po
        import svtme_b200  # noqa: F401
        from svt_av1_psy_b200 import abi
        self.wl, self.frames, self.threads, self.po, self.abi = wl, frames, threads, po, abi
        self.use_ref = po.ref_available()
        self.simd = 2 if "avx512bw" in open("/proc/cpuinfo").read() else 1
        self.pool = ThreadPoolExecutor(max_workers=threads)
        self.cached = {}  # frame index -> last step that used it
        self.step_no = 0
        self.table_seen = None
        if self.use_ref:
            self.lib = po.ref_lib()
            self.lib.ref_plane_cache_build.argtypes = [C.c_void_p, C.c_int, C.c_int]
            self.lib.ref_plane_cache_drop.argtypes = [C.c_void_p]
            self.lib.ref_plane_cache_enable(1)
        else:
            po.oracle_lib()

    @property
    def kind(self):
        return "reference" if self.use_ref else "port"

    @property
    def kernels(self):
        if not self.use_ref:
            return "oracle C port"
        return "reference %s ME kernels" % ("AVX-512 SAD loop + AVX2" if self.simd == 2 else "AVX2")

    def _one(self, poc):
        wl, po, frames = self.wl, self.po, self.frames
        plan = wl.plan(poc)
        src = frames[poc % wl.n_frames]
        refs = [[frames[p % wl.n_frames] for p in plan["l0"]], [frames[p % wl.n_frames] for p in plan["l1"]]]
        if self.use_ref:
            pin = po.make_pic_in(src, refs, [plan["l0"], plan["l1"]], poc, wl.preset, qp=wl.qp, temporal_layer=plan["tl"],
                                 hier_levels=wl.hier, is_ref=plan["is_ref"], simd=self.simd, mrp=wl.mrp)
            res = self.abi.MePicResults()  # no dump wanted: the stage writes its results into the PCS arrays as in the encoder
            rc = self.lib.ref_me_picture(C.byref(pin), C.byref(res), None, None)
            assert rc == 0
        else:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from util import orc_me, params_from_encoder_inputs
            prm = params_from_encoder_inputs(src, refs, [plan["l0"], plan["l1"]], poc, wl.preset, qp=wl.qp,
                                             temporal_layer=plan["tl"], hier_levels=wl.hier, is_ref=plan["is_ref"], mrp=wl.mrp)
            orc_me(src, refs, prm, debug=False)

    def _build(self, f):
        fr = self.frames[f]
        self.lib.ref_plane_cache_build(fr.ctypes.data, self.wl.w, self.wl.h)

    def step(self, pocs):
        """ME stage of `pocs` -> wall seconds (pyramids of pictures not seen yet are built inside the timed region)"""
        wl = self.wl
        need = set()
        for poc in pocs:
            plan = wl.plan(poc)
            need.update(p % wl.n_frames for p in [poc] + plan["l0"] + plan["l1"])
        t0 = time.time()
        if self.use_ref:
            list(self.pool.map(self._build, [f for f in sorted(need) if f not in self.cached]))
        list(self.pool.map(self._one, pocs))
        sec = time.time() - t0
        if self.use_ref:
            self.table_seen = self.lib.ref_rtcd_table_at_last_b64()
            assert self.table_seen == self.simd, "the reference ran table %s, not %s" % (self.table_seen, self.simd)
            for f in need:
                self.cached[f] = self.step_no
            for f in [f for f, last in self.cached.items() if last < self.step_no - 6]:  # outside any reference reach
                self.lib.ref_plane_cache_drop(self.frames[f].ctypes.data)
                del self.cached[f]
        self.step_no += 1
        return sec

    def close(self):
        self.pool.shutdown()
        if self.use_ref:
            self.lib.ref_plane_cache_drop(None)
            self.lib.ref_plane_cache_enable(0)


def cpu_baseline(wl, frames, budget_s=12.0):
    """bounded sample: whole mini-GOPs of the same workload until ~budget_s of wall time (at least one after a warm-up)"""
    cores = os.cpu_count() or 1
    arm = CpuArm(wl, frames, cores)
    arm.step(wl.order(wl.mgop))  # warm-up mini-GOP (also fills the plane cache with the reference pictures)
    n_pics, sec, m = 0, 0.0, 2
    while sec < budget_s and m < 2 + 256:
        pocs = wl.order(m * wl.mgop)
        sec += arm.step(pocs)
        n_pics += len(pocs)
        m += 1
    out = {"value": round(n_pics / sec, 3), "unit": "frames/s", "cores": cores, "kind": arm.kind,
           "sample": "%d mini-GOPs (%d pictures) of the same workload after one warm-up mini-GOP, one picture per host thread, %s "
                     "(table probed at the b64 call: %s), pyramid of each picture built once inside the timed region"
                     % (m - 2, n_pics, arm.kernels, arm.table_seen),
           "seconds": round(sec, 2)}
    arm.close()
    return out


def run_reference(a):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    wl = Workload(a)
    frames = make_clip(wl, 0)
    cores = os.cpu_count() or 1
    arm = CpuArm(wl, frames, cores)
    secs = []
    for s in range(a.warmup + a.steps):
        sec = arm.step(wl.order((1 + s) * wl.mgop))  # one step = one mini-GOP, as in our arm
        if s >= a.warmup:
            secs.append(sec)
    total = sum(secs)
    value = wl.mgop * a.steps / total
    line = {
        "impl": "reference", "metric": "open-loop ME frames/s", "value": round(value, 3), "unit": "frames/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(1e3 * total / a.steps, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl.name, "pictures_per_step": wl.mgop},
        "cpu_baseline": {"value": round(value, 3), "unit": "frames/s", "cores": cores, "kind": arm.kind,
                         "sample": "one mini-GOP (%d pictures) per step, one picture per host thread, %s (table probed at the b64 "
                                   "call: %s)" % (wl.mgop, arm.kernels, arm.table_seen)},
        "e2e": {"value": round(value, 3), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    arm.close()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    args = parse_args()
    import svtme_b200  # noqa: F401,E402
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
