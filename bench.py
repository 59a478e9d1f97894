#!/usr/bin/env python
"""bench.py — images/s of the detection-head hot path (match+loss and decode+NMS) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic head outputs of the
BASELINE.json shape (default: configs[1], EfficientDet-D0 512x512, batch 64, 90 classes,
<=100 GT boxes/image, 9 anchors/cell = 49 104 anchors/image):
    training side   FocalLoss forward + backward (assignment, focal + smooth-L1, gradients)
    inference side  decode + clip + score threshold + sort + NMS for every image of the batch
`value` = images/s with inputs resident in HBM (device-timed, max over ranks);
`e2e`   = the same step from PINNED HOST buffers through the public module API, host->device
          copies of the inputs and device->host reads of the results inside the timed region;
`roofline` = the dominant kernel (k7_loss, forward+backward) against the measured HBM peak;
`cpu_baseline` = the oracle port (oracle/restate.py, torch-CPU, all host threads) on a
          bounded sample of the same workload (rank 0, N=1 only).
N > 1 (torchrun): every rank owns a full batch (weak scaling; `--shard` splits the workload's
batch over the ranks instead: strong scaling, BASELINE configs[3] and [4]); the four loss sums are
exchanged inside k7_loss over NVLink peer memory (`--exchange nccl`: one 16-byte NCCL all-reduce).
`--impl reference` times the reference's CPU algorithm for the same step (the oracle port;
the reference is pure Python/PyTorch, there is nothing to compile) and prints the same line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "images/s for match+loss and decode+NMS"
UNIT = "images/s"
DEFAULT_WORKLOAD = "c2_effdet_d0_512"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--batch", type=int, default=0, help="override the per-GPU batch (debug)")
    ap.add_argument("--shard", action="store_true",
                    help="strong scaling: split the workload's batch over the ranks (BASELINE configs[3], [4]) "
                         "instead of one full batch per rank")
    ap.add_argument("--exchange", default="auto", choices=["auto", "nccl"],
                    help="N>1: how the four loss sums are combined: fused into k7_loss over peer memory (auto; "
                         "falls back to NCCL if the mailboxes cannot be mapped) or NCCL all_reduce")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip parts.configs (stage times of the other BASELINE configs)")
    ap.add_argument("--no-shard-part", action="store_true", help="skip parts.c5_shard (EfficientDet-D7 batch 64 split over the ranks)")
    ap.add_argument("--serial", action="store_true", help="run both halves of the step on one stream")
    ap.add_argument("--eager", action="store_true", help="launch the timed steps eagerly instead of replaying the captured graph")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def ncu_traffic(workload, batch):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed
    `ncu --set full` capture of this workload (profiles/traffic.json); None if there is none."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            t = json.load(f)
        e = t.get(workload)
        if e and e.get("batch") == batch:
            return e["k7_loss_pipe_kernel_grad_bytes"]
    except Exception:
        pass
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------ CPU arm (oracle port)
def cpu_step(R, cls, reg, anc, ann, hw):
    """The reference's algorithm for one step on CPU tensors: FocalLoss fwd+bwd, then the eval tail."""
    c = cls.detach().requires_grad_(True)
    r = reg.detach().requires_grad_(True)
    out = R.focal_loss(c, r, anc, ann)
    (out.cls_mean + out.reg_mean).sum().backward()
    with torch.no_grad():
        R.detect_batch(cls, reg, anc, hw[0], hw[1])
    return out


def cpu_inputs(wl, n_images, seed=0):
    import b200det
    from oracle import restate as R
    cls, reg, ann = b200det.synth.make_inputs(wl, seed=seed, device="cpu", batch=n_images)
    anc = torch.from_numpy(R.retina_anchors(wl.height, wl.width, **wl.anchor_kwargs()))[None]
    return R, cls, reg, ann, anc


def cpu_calibrate(wl, hw, n0=4):
    """Seconds per image of the CPU arm, measured on a warm pass over n0 images."""
    n0 = min(n0, wl.batch)
    R, cls, reg, ann, anc = cpu_inputs(wl, n0)
    cpu_step(R, cls, reg, anc, ann, hw)                       # warm-up (allocator, thread pool)
    t0 = time.perf_counter()
    cpu_step(R, cls, reg, anc, ann, hw)
    return (time.perf_counter() - t0) / n0


def cpu_baseline(wl, budget_s=12.0):
    """Oracle port timed on the host cores on a bounded sample of the workload."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hw = (wl.height, wl.width)
    per_img = cpu_calibrate(wl, hw)
    n = int(max(1, min(wl.batch, budget_s / max(per_img, 1e-6))))
    R, cls, reg, ann, anc = cpu_inputs(wl, n)
    cpu_step(R, cls, reg, anc, ann, hw)                       # warm pass at this size
    t0 = time.perf_counter()
    cpu_step(R, cls, reg, anc, ann, hw)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} of {wl.batch} images of {wl.name}, 1 timed pass after 1 warm pass (FocalLoss fwd+bwd + "
                      f"decode/NMS), torch-CPU oracle port of the reference, {cores} threads"}


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hw = (wl.height, wl.width)
    per_img = cpu_calibrate(wl, hw)
    total_steps = args.steps + args.warmup
    # bound the whole run to ~2 minutes (large batches run slower per image: 2x margin)
    n = int(max(1, min(wl.batch, 60.0 / max(per_img * total_steps, 1e-6))))
    R, cls, reg, ann, anc = cpu_inputs(wl, n)
    for _ in range(args.warmup):
        cpu_step(R, cls, reg, anc, ann, hw)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(R, cls, reg, anc, ann, hw)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = (f"{n} of {wl.batch} images of {wl.name} per step; oracle port (torch-CPU restatement of the "
              f"reference, same ATen ops), {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(wl, wl.batch, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(wl, batch, gpus):
    """The `config` object: identical in both arms (b200 / reference) for the same command line."""
    return {"workload": wl.name, "image": [wl.height, wl.width], "batch_per_gpu": batch, "global_batch": batch * gpus,
            "classes": wl.classes, "anchors_per_image": wl.anchors, "max_gt_per_image": wl.max_gt,
            "step": "FocalLoss fwd+bwd + decode/clip/threshold/sort/NMS on every image",
            "cache": (f"inputs ({batch * wl.anchors * (4 * wl.classes + 16) / 1e9:.2f} GB per rank and pass) are larger than "
                      "the 126 MB L2; no flush needed") if batch * wl.anchors * (4 * wl.classes + 16) > 4 * 126e6 else
                     "inputs fit (partly) in the 126 MB L2: L2-warm numbers, not comparable with the HBM roofline",
            "parallelism": f"image-sharded x{gpus}, the four loss sums exchanged once per step" if gpus > 1 else "single GPU"}


# ------------------------------------------------------------------------------ B200 arm
def timed(fn, steps, warmup, stream_sync=True):
    """Average milliseconds per call over `steps` calls, CUDA events on the current stream."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def bind_to_gpu_numa(local, local_world):
    """Bind this process (and with it the first-touch placement of its pinned host buffers) to a disjoint share of
    the cores of the NUMA node its GPU hangs on.  Returns a description for the JSON line."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = -1
        try:
            with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
                node = int(f.read().strip())
        except Exception:
            pass
        allowed = sorted(os.sched_getaffinity(0))
        cores = allowed
        if node >= 0:
            with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                lst = []
                for part in f.read().strip().split(","):
                    a, _, b = part.partition("-")
                    lst += list(range(int(a), int(b or a) + 1))
            cores = [c for c in lst if c in set(allowed)] or allowed
        # the ranks whose GPUs share the node split its cores; without topology knowledge: an even split by local rank
        per = max(1, len(cores) // max(1, local_world))
        mine = cores[(local * per) % len(cores):][:per] or cores
        os.sched_setaffinity(0, mine)
        torch.set_num_threads(max(1, min(len(mine), 8)))
        return {"gpu_pci": bdf, "numa_node": node, "cores": f"{mine[0]}-{mine[-1]} ({len(mine)})"}
    except Exception as e:
        return {"error": str(e)[:120]}


class HotPath:
    """One workload on one GPU: inputs, modules and the step of the bench."""

    def __init__(self, wl, B, dev, seed, world=1, peer="auto", spare_sms=None):
        import b200det
        self.b200det, self.wl, self.B, self.dev = b200det, wl, B, dev
        self.world, self.no_defer = world, bool(os.environ.get("K7_BENCH_NO_DEFER"))
        self.hw = (wl.height, wl.width)
        self.cls, self.reg, self.ann = b200det.synth.make_inputs(wl, seed=seed, device=dev, batch=B)
        self.img = torch.empty(B, 3, wl.height, wl.width, device=dev)   # only its shape is read (anchors.py:40)
        self.anchors_mod = b200det.Anchors(**wl.anchor_kwargs())
        # large images: the NMS kernels (1024-thread CTAs) cannot share an SM with the loss stage's streaming CTAs,
        # so the stream leaves them a few SMs (it stays HBM-bound) and both run side by side in step()
        # (measured on D7, two-stream step: 8 images 0.763 -> 0.685 ms with 24, 16 images 1.364 -> 1.284 with 12,
        # 32 images 2.585 -> 2.465 with 16; at 64 images the gain is inside the run-to-run spread, and small images
        # (C2) lose: their NMS tail is short and their loss stream needs every SM)
        large = wl.anchors >= 262144
        self.spare_sms = ((24 if B < 12 else 16 if B <= 32 else 0) if large else 0) if spare_sms is None else spare_sms
        self.loss_mod = b200det.FocalLoss(sharded=world > 1, peer_reduce=peer, spare_sms=self.spare_sms)
        self.cls_g = self.cls.detach().requires_grad_(True)
        self.reg_g = self.reg.detach().requires_grad_(True)
        self.side = torch.cuda.Stream(device=dev)

    def train_side(self):
        self.cls_g.grad = None
        self.reg_g.grad = None
        lc, lr = self.loss_mod(self.cls_g, self.reg_g, self.anchors_mod(self.img), self.ann)
        (lc + lr).sum().backward()
        return lc, lr

    def train_start(self):
        """The training side in its two-phase form (FocalLoss.start): gradients now, the global loss values when
        .losses() is called - nothing on this stream waits for a peer."""
        self.cls_g.grad = None
        self.reg_g.grad = None
        pend = self.loss_mod.start(self.cls_g, self.reg_g, self.anchors_mod(self.img), self.ann)
        pend.backward()
        return pend

    def fwd_only(self):
        with torch.no_grad():
            return self.loss_mod(self.cls, self.reg, self.anchors_mod(self.img), self.ann)

    def eval_side(self, sync=False):
        return self.b200det.detect(self.cls, self.reg, self.anchors_mod(self.img), self.hw, sync=sync)

    def step_serial(self):
        self.train_side()
        self.eval_side()

    def step(self):
        """The two halves of a step are independent: decode/NMS runs on a second stream, so its latency-bound tail
        (sort / pair tests / resolve: SM-light) overlaps the training side's streaming kernels and vice versa.
        Workspaces are per stream (losses._Workspace), results are joined before the step ends."""
        cur = torch.cuda.current_stream(self.dev)
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            self.eval_side()
        if self.world > 1 and not self.no_defer:
            pend = self.train_start()
            cur.wait_stream(self.side)
            return pend.losses()                 # the exchange is joined at the very end of the step
        self.train_side()
        cur.wait_stream(self.side)

    def graphed(self, fn=None):
        """The step captured once in a CUDA graph (b200det.GraphedStep) - None when the capture fails."""
        try:
            return self.b200det.GraphedStep(fn or self.step, warmup=3, device=self.dev)
        except Exception as e:                                   # never let the capture break the line
            self.graph_error = str(e)[:200]
            torch.cuda.synchronize()
            return None

    def bytes(self, surv=0):
        wl, B = self.wl, self.B
        A, C, M = wl.anchors, wl.classes, wl.max_gt
        return (B * (A * (4 * C + 16) + 20 * M), B * (A * (8 * C + 32) + 20 * M), B * A * (4 * C + 16) + 32 * surv)

    def stage_times(self, k, peak):
        """Stage times of this workload (ms) and their fraction of the HBM peak (SURVEY 8d bytes)."""
        from b200det import losses as BL
        A = self.wl.anchors
        anc = self.anchors_mod(self.img)
        sp = self.b200det.spec_of(anc)
        self.loss_mod.cfg["spare_sms"] = 0            # the stages by themselves: the loss stream on all SMs
        ms_fwd = timed(self.fwd_only, k, 3)
        ms_fb = timed(self.train_side, k, 3)
        ms_det = timed(self.eval_side, k, 3)
        g_train = self.graphed(self.train_side)
        ms_graph_fb = timed(g_train, k, 3) if g_train is not None else None
        del g_train
        self.loss_mod.cfg["spare_sms"] = self.spare_sms   # the step: both halves side by side
        ms_step = timed(self.step, k, 3)
        g_step = self.graphed()
        ms_graph = timed(g_step, k, 3) if g_step is not None else None
        del g_step
        ms_match = timed(lambda: BL.match_scatter(sp, self.ann, A, 0.35, 0.45), k, 3)
        surv = int(self.eval_side(sync=True).survivors.sum())
        bf, bfb, bd = self.bytes(surv)
        g = lambda nbytes, ms: nbytes / (ms * 1e-3) / 1e9 / peak
        return {"batch": self.B, "anchors_per_image": A, "classes": self.wl.classes,
                "k7_match_ms": ms_match, "match_loss_fwd_ms": ms_fwd, "match_loss_fwd_hbm_frac": g(bf, ms_fwd),
                "match_loss_fwd_bwd_ms": ms_fb, "match_loss_fwd_bwd_hbm_frac": g(bfb, ms_fb),
                "decode_nms_ms": ms_det, "decode_nms_hbm_frac": g(bd, ms_det), "survivors_per_image": surv / self.B,
                "step_ms": ms_step, "step_graph_ms": ms_graph,
                "match_loss_fwd_bwd_graph_ms": ms_graph_fb,
                "match_loss_fwd_bwd_graph_hbm_frac": g(bfb, ms_graph_fb) if ms_graph_fb else None,
                "images_per_s": self.B / (ms_graph or ms_step) * 1e3}


def run_b200(args, wl):
    import torch.distributed as dist
    import b200det
    from b200det import _lib as L
    from b200det import losses as BL

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (b200det has no CPU path); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    binding = bind_to_gpu_numa(local, local_world) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = L.lib()                                              # fail loudly if libk7det.so is missing

    B = args.batch or wl.batch
    if args.shard:
        if B % world:
            raise SystemExit(f"--shard: batch {B} does not divide over {world} ranks")
        B //= world                                            # rank r owns images [r*B, (r+1)*B) of the global batch
    A, C, M = wl.anchors, wl.classes, wl.max_gt
    peer = "auto" if args.exchange == "auto" else False
    hp = HotPath(wl, B, dev, seed=rank, world=world, peer=peer)
    cls, reg, ann, img, hw = hp.cls, hp.reg, hp.ann, hp.img, hp.hw
    anchors_mod, loss_mod = hp.anchors_mod, hp.loss_mod
    step_eager = hp.step_serial if args.serial else hp.step
    graphed = None if args.eager else hp.graphed(step_eager)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_ranks(fn, steps, warmup):
        """ms per call, barrier + synchronize on both sides, max over ranks."""
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    def pick(g, eager):
        """Graph replay or eager launches, whichever is faster for this step on these ranks (max over ranks): replay
        removes the host from the loop, but the graph executor interleaves the two branches of a two-stream step in
        its own order, which at few images per GPU can be the worse one."""
        ok = torch.tensor([1 if g is not None else 0], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if not int(ok.item()):
            return eager, False
        tg, te = timed_ranks(g, 5, 3), timed_ranks(eager, 5, 3)
        return (g, True) if tg <= te else (eager, False)

    step, use_graph = pick(graphed, step_eager)
    if not use_graph:
        graphed = None

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()

    # ---- value: K steps, device resident, barrier + synchronize on both sides, max over ranks
    W = max(args.warmup, 3)
    for _ in range(W):
        step()
    barrier()
    n0 = lib.k7_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = int(lib.k7_launch_count() - n0)
    if graphed is not None:
        launches = graphed.launches * args.steps                # a replay makes no API call: the graph holds them
    ms_total = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = float(ms_total.item()) / args.steps
    value = world * B / (ms_step * 1e-3)

    peak, peak_src = peaks()
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if args.shard else "weak",
        "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(wl, B, world),
        "gpu_launches": launches,
        "gpu_launches_note": "counted by libk7det (k7_launch_count): every kernel of the library launched in the timed region "
                             "(graph replay: the kernels recorded in the captured step x steps); autograd adds a few tiny "
                             "ATen kernels per step, memsets are not kernels",
        "step_schedule": "serial (one stream)" if args.serial else
                         "decode/NMS on a second CUDA stream next to FocalLoss fwd+bwd, joined at the end of the step",
        "step_launch": "CUDA graph replay of the step through the public modules (b200det.GraphedStep); static input tensors, "
                       "all work redone every replay" if graphed is not None else "eager",
    }
    if graphed is None and not args.eager:
        out["graph_error"] = getattr(hp, "graph_error", None)
    if world > 1:
        out["loss_exchange"] = loss_mod.exchange
        out["host_binding"] = binding

    # ---- exchange check (N > 1): the fused sums of one call against an all-reduce of the ranks' own sums
    if world > 1:
        hp.train_side()
        fused = loss_mod.last_sums.detach().clone()
        local_mod = b200det.FocalLoss()
        with torch.no_grad():
            local_mod(cls, reg, anchors_mod(img), ann)
        ref = local_mod.last_sums.detach().clone()
        dist.all_reduce(ref, op=dist.ReduceOp.SUM)
        ok = bool(torch.isfinite(fused).all()) and bool(torch.allclose(fused, ref, rtol=1e-6, atol=0.0))
        okt = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        out["exchange_check"] = bool(okt.item())
        out["exchange_sums"] = {"fused": [float(x) for x in fused.tolist()], "all_reduce_of_local": [float(x) for x in ref.tolist()]}

    # ---- parts and the roofline of the dominant kernel (every rank runs them; rank 0 reports)
    k = max(args.steps, 10)
    ms_fwd = timed(hp.fwd_only, k, 3)
    ms_fb = timed(hp.train_side, k, 3)
    ms_det = timed(hp.eval_side, k, 3)
    ms_serial = timed(hp.step_serial, k, 3)
    ms_eager = timed(hp.step, k, 3)
    replayed = {}
    for name, fn, nbytes in (("match_loss_fwd", hp.fwd_only, None), ("match_loss_fwd_bwd", hp.train_side, None),
                             ("decode_nms", hp.eval_side, None)):
        g_ = hp.graphed(fn)
        if g_ is not None:
            replayed[name] = timed(g_, k, 3)
        del g_
    anc = anchors_mod(img)
    sp_ = b200det.spec_of(anc)
    kw = dict(alpha=0.25, beta=1.0 / 9.0, target_div=(0.1, 0.1, 0.2, 0.2), yx_targets=False, mean_divisor=B)
    dreg0 = torch.empty_like(reg)
    ms_match = timed(lambda: BL.match_scatter(sp_, ann, A, 0.35, 0.45, zero_fill=dreg0), k, 3)
    arr_, gen_, consume_ = BL.match_scatter(sp_, ann, A, 0.35, 0.45, zero_fill=dreg0)
    best_ = (arr_, gen_, consume_, 0.35, 0.45)
    ms_k_fb = timed(lambda: BL.loss_stage(anc, cls, reg, ann, None, None, want_grad=True, best=best_, dreg_zeroed=dreg0, **kw), k, 3)
    ms_k_f = timed(lambda: BL.loss_stage(anc, cls, reg, ann, None, None, want_grad=False, best=best_, **kw), k, 3)
    surv = int(hp.eval_side(sync=True).survivors.sum())
    bytes_fwd, bytes_fb, bytes_det = hp.bytes(surv)

    def gbs(nbytes, ms):
        return nbytes / (ms * 1e-3) / 1e9

    out["parts"] = {
        "match_loss_fwd": {"ms": ms_fwd, "images_per_s": B / ms_fwd * 1e3, "hbm_frac": gbs(bytes_fwd, ms_fwd) / peak},
        "match_loss_fwd_bwd": {"ms": ms_fb, "images_per_s": B / ms_fb * 1e3, "hbm_frac": gbs(bytes_fb, ms_fb) / peak},
        "decode_nms": {"ms": ms_det, "images_per_s": B / ms_det * 1e3, "hbm_frac": gbs(bytes_det, ms_det) / peak,
                       "survivors_per_image": surv / B},
        "step_eager": {"ms": ms_eager, "images_per_s": B / ms_eager * 1e3,
                       "note": "the two-stream step launched eagerly from Python (the host sets the pace at small batches)"},
        "step_one_stream": {"ms": ms_serial, "images_per_s": B / ms_serial * 1e3,
                            "note": "the same step, eager, with both halves on one stream (what the previous round reported)"},
        "k7_match_ms": ms_match, "k7_loss_fwd_ms": ms_k_f, "k7_loss_fwd_bwd_ms": ms_k_fb,
        "k7_loss_fwd_hbm_frac": gbs(bytes_fwd, ms_k_f) / peak,
    }
    # the same three module calls replayed from a CUDA graph: the kernels without the host's launch gaps
    rb = {"match_loss_fwd": bytes_fwd, "match_loss_fwd_bwd": bytes_fb, "decode_nms": bytes_det}
    out["parts"]["replayed"] = {n: {"ms": t, "images_per_s": B / t * 1e3, "hbm_frac": gbs(rb[n], t) / peak}
                                for n, t in replayed.items()}
    del dreg0, arr_, best_
    if world == 1:
        g1 = hp.graphed(hp.step_serial)
        if g1 is not None:
            ms_graph = timed(g1, k, 3)
            out["parts"]["cuda_graph_step_one_stream"] = {"ms": ms_graph, "images_per_s": B / ms_graph * 1e3,
                                                         "note": "graph replay of the ONE-stream step: what the second stream buys is "
                                                                 "the difference to the headline"}
        del g1
    # the same stages fed with the head's pre-sigmoid logits (sigmoid fused), next to the
    # stock elementwise sigmoid pass (and its autograd backward) that they make unnecessary
    z = torch.logit(cls)
    z_g = z.detach().requires_grad_(True)
    reg_g = hp.reg_g
    loss_lg = b200det.FocalLoss(from_logits=True)

    def train_side_logits():
        z_g.grad = None
        reg_g.grad = None
        lc, lr = loss_lg(z_g, reg_g, anchors_mod(img), ann)
        (lc + lr).sum().backward()

    ms_fb_lg = timed(train_side_logits, k, 3)
    ms_det_lg = timed(lambda: b200det.detect(z, reg, anchors_mod(img), hw, sync=False, from_logits=True), k, 3)
    ms_sig = timed(lambda: torch.sigmoid(z), k, 3)
    out["parts"]["from_logits"] = {
        "match_loss_fwd_bwd_ms": ms_fb_lg, "decode_nms_ms": ms_det_lg,
        "torch_sigmoid_pass_ms_saved_per_side": ms_sig,
        "note": "FocalLoss(from_logits=True) / detect(from_logits=True): the head's sigmoid runs in registers"}
    # ... and with the conv heads' own per-level NCHW outputs (no permute/contiguous/view/cat either)
    sp = b200det.spec_of(anchors_mod(img))
    Kc = sp.cell_anchors
    lv = [z[:, sp.offset[l]:sp.offset[l + 1], :].reshape(B, sp.fh[l], sp.fw[l], Kc * C).permute(0, 3, 1, 2).contiguous()
          for l in range(sp.num_levels)]
    lv_g = [t.detach().requires_grad_(True) for t in lv]

    def train_side_levels():
        for t in lv_g:
            t.grad = None
        reg_g.grad = None
        lc, lr = loss_lg(lv_g, reg_g, anchors_mod(img), ann)
        (lc + lr).sum().backward()

    ms_fb_lv = timed(train_side_levels, k, 3)
    ms_det_lv = timed(lambda: b200det.detect(lv, reg, anchors_mod(img), hw, sync=False, from_logits=True), k, 3)
    ms_layout = timed(lambda: torch.cat([t.permute(0, 2, 3, 1).contiguous().view(B, -1, C) for t in lv], dim=1), k, 3)
    out["parts"]["head_levels"] = {
        "match_loss_fwd_bwd_ms": ms_fb_lv, "decode_nms_ms": ms_det_lv,
        "torch_layout_pass_ms_saved_per_side": ms_layout,
        "note": "FocalLoss / detect on the list of per-level [B, K*C, fh, fw] logits (retinanet.py:159-178 layout code skipped)"}
    for t in lv_g:
        t.grad = None
    z_g.grad = None
    del z, z_g, lv, lv_g
    # the evaluators' per-image host loops (coco_eval.py:25-57) as one device call
    det_e = hp.eval_side(sync=True)
    scales_e = [1.0 + 0.01 * i for i in range(B)]
    ms_pack = timed(lambda: b200det.eval_pack(det_e, scales_e, mode="coco", num_classes=C), k, 3)
    out["parts"]["eval_tail"] = {"eval_pack_coco_ms": ms_pack, "detections": int(det_e.count.sum()),
                                 "note": "boxes/scale, xywh, score cut, packed float64 [N,7] rows for the batch (k7_eval_pack, "
                                         "3 launches + 1 sync); the host loop it replaces is timed in cpu_baseline.eval_tail_ms"}
    eval_inputs = (det_e.scores.cpu().numpy(), det_e.labels.cpu().numpy(), det_e.boxes.cpu().numpy(),
                   det_e.count.cpu().numpy(), scales_e, list(range(B)), list(range(C)))
    del det_e
    # the dominant kernel: the k7_loss call with gradients (list + stream + serve kernels).  Its bytes: classifications
    # in, their gradient out (8 C per anchor) + the annotations; the 16 B/anchor zero fill of d(reg) is k7_match_scatter's
    # and reg / d(reg) are touched for the ~3 % positive rows only, so neither is counted here.
    bytes_loss = B * (A * 8 * C + 20 * M)
    ach = gbs(bytes_loss, ms_k_fb)
    traffic = ncu_traffic(wl.name, B)
    out["roofline"] = {"bound": "hbm", "kernel": "k7_loss_pipe_kernel<grad> (timed as the k7_loss call: list kernel + stream kernel + serve/reduction kernel)",
                       "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                       "traffic": traffic,
                       "traffic_source": "committed ncu --set full capture (profiles/traffic.json), not measured in this run" if traffic else None,
                       "algorithmic_bytes_per_launch": bytes_loss, "avg_launch_ms": ms_k_fb, "peak_source": peak_src}

    # ---- e2e: pinned host inputs -> H2D -> public modules -> D2H of the results, every step
    if not args.no_e2e:
        h_cls, h_reg, h_ann = (t.cpu().pin_memory() for t in (cls, reg, ann))
        d_cls, d_reg, d_ann = torch.empty_like(cls), torch.empty_like(reg), torch.empty_like(ann)
        d2h = [0]

        # The copy of the big tensor runs on its own stream in image chunks; decode/NMS (per-image work, the public
        # detect() on a slice) of a chunk starts as soon as the chunk has landed, so only the last chunk's detect
        # and the batch-wide FocalLoss are left when the copy ends.  Everything still goes through the modules.
        copy_stream = torch.cuda.Stream()
        n_chunks = 8 if B % 8 == 0 and B >= 16 else 1
        per = B // n_chunks

        def e2e_step():
            cur = torch.cuda.current_stream()
            copy_stream.wait_stream(cur)                      # the previous step has consumed d_cls / d_reg / d_ann
            arrived = []
            with torch.cuda.stream(copy_stream):
                d_reg.copy_(h_reg, non_blocking=True)
                d_ann.copy_(h_ann, non_blocking=True)
                for i in range(n_chunks):
                    d_cls[i * per:(i + 1) * per].copy_(h_cls[i * per:(i + 1) * per], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(copy_stream)
                    arrived.append(ev)
            dets = []
            for i in range(n_chunks):
                cur.wait_event(arrived[i])
                sl = slice(i * per, (i + 1) * per)
                dets.append(b200det.detect(d_cls[sl], d_reg[sl], anchors_mod(img[sl]), hw, sync=False))
            c = d_cls.detach().requires_grad_(True)
            r = d_reg.detach().requires_grad_(True)
            lc, lr = loss_mod(c, r, anchors_mod(img), d_ann)
            (lc + lr).sum().backward()
            losses = torch.cat([lc.detach(), lr.detach()]).cpu()
            counts = torch.cat([d.count for d in dets]).cpu()
            kmax = max(int(counts.max()), 1)
            assert kmax <= dets[0].scores.shape[1], "detect(sync=False) returned fewer rows than an image kept: pass cap="
            res = (torch.cat([d.scores[:, :kmax] for d in dets]).cpu(), torch.cat([d.labels[:, :kmax] for d in dets]).cpu(),
                   torch.cat([d.boxes[:, :kmax] for d in dets]).cpu())
            d2h[0] = losses.numel() * 4 + B * 2 * 4 + sum(t.numel() * t.element_size() for t in res)

        n_e2e = max(3, min(args.steps, 10))
        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        torch.cuda.synchronize()
        ms_e2e = torch.tensor([(time.perf_counter() - t0) * 1e3 / n_e2e], device=dev)
        if world > 1:
            dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
        h2d = sum(t.numel() * 4 for t in (cls, reg, ann))
        out["e2e"] = {"value": world * B / (float(ms_e2e.item()) * 1e-3), "unit": UNIT,
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h[0],
                      "ms_per_step": float(ms_e2e.item()), "steps": n_e2e,
                      "h2d_gb_per_s_all_ranks": world * h2d / (float(ms_e2e.item()) * 1e-3) / 1e9,
                      "note": "pinned host -> device copies (chunked on a copy stream, detect() of a chunk overlaps the next copy) + "
                              "FocalLoss fwd+bwd + detect + results to host; PCIe-bound: the copy alone is 20.4 ms at 55 GB/s per GPU"}
        del h_cls, h_reg, h_ann, d_cls, d_reg, d_ann

    # ---- the other BASELINE configurations on one GPU, at their BASELINE batch (stage times, not bench lines)
    del graphed, step, step_eager
    del hp, cls, reg, ann, img, anchors_mod, loss_mod, loss_lg, reg_g
    torch.cuda.empty_cache()
    if world == 1 and not args.no_configs and not args.batch and not args.shard:
        out["parts"]["configs"] = {}
        for name in ("c1_retinanet_r50_512", "c3_ksevendet_thermal_512x640", "c4_effdet_d3_896", "c5_effdet_d7_1536"):
            if name == wl.name:
                continue
            try:
                w2 = b200det.synth.WORKLOADS[name]
                h2 = HotPath(w2, w2.batch, dev, seed=0)
                out["parts"]["configs"][name] = h2.stage_times(10, peak)
                del h2
            except Exception as e:
                out["parts"]["configs"][name] = {"error": str(e)[:200]}
            torch.cuda.empty_cache()

    # ---- north_star's multi-GPU target: EfficientDet-D7 1536^2, batch 64 SPLIT over the ranks (strong scaling),
    # with the same run's one-GPU time of the whole batch next to it
    if not args.no_shard_part and not args.batch and not args.shard:
        try:
            w5 = b200det.synth.WORKLOADS["c5_effdet_d7_1536"]
            if w5.batch % world == 0:
                Bs = w5.batch // world
                h5 = HotPath(w5, Bs, dev, seed=rank, world=world, peer=peer)
                g5 = None if args.eager else h5.graphed()
                step5, graph5 = pick(g5, h5.step)
                ms_n = timed_ranks(step5, 10, 3)
                del g5, step5
                rec = {"workload": w5.name, "global_batch": w5.batch, "images_per_gpu": Bs, "n_gpus": world,
                       "ms_per_step": ms_n, "images_per_s": w5.batch / ms_n * 1e3, "scaling": "strong",
                       "step_launch": "CUDA graph replay" if graph5 else "eager",
                       "spare_sms": h5.spare_sms}
                if world > 1:
                    rec["loss_exchange"] = h5.loss_mod.exchange
                del h5
                torch.cuda.empty_cache()
                hs = HotPath(w5, Bs, dev, seed=rank)             # the rank's shard without the exchange: stage times
                st = hs.stage_times(10, peak)
                del hs
                torch.cuda.empty_cache()
                rec["stages_rank0"] = st
                ms_1 = ms_n
                if world > 1:
                    dist.barrier()
                    if rank == 0:                                   # the whole batch on ONE GPU, same run, same code
                        h1 = HotPath(w5, w5.batch, dev, seed=0)
                        g1 = None if args.eager else h1.graphed()
                        ms_1 = timed(h1.step, 5, 2)
                        if g1 is not None:
                            ms_1 = min(ms_1, timed(g1, 5, 2))      # the better of the two, as for the sharded step
                        del g1
                        del h1
                        torch.cuda.empty_cache()
                    dist.barrier()
                rec["one_gpu_ms_per_step"] = ms_1
                rec["speedup_vs_one_gpu"] = ms_1 / ms_n
                rec["efficiency"] = ms_1 / ms_n / world
                out["parts"]["c5_shard"] = rec
        except Exception as e:
            out["parts"]["c5_shard"] = {"error": str(e)[:300]}
            if world > 1:
                raise
        torch.cuda.empty_cache()

    if rank == 0:
        out["clocks"] = sampler.stop()
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(wl)
            from oracle import restate as R_
            t0 = time.perf_counter()
            R_.eval_rows_coco(*eval_inputs)
            out["cpu_baseline"]["eval_tail_ms"] = (time.perf_counter() - t0) * 1e3
        print(json.dumps(out))
    bad = world > 1 and not out.get("exchange_check", True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if bad:
        raise SystemExit("bench.py: the fused loss-sum exchange disagrees with the all-reduce of the ranks' sums")


def main():
    args = parse()
    import b200det
    wl = b200det.synth.WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()
