"""Reference arm of bench.py (``bench.py --impl reference``).

Runs the headline config (BASELINE.json config 3: DeepFM, 125 M keys per GPU, dim 16, 26 slots,
batch 4096 per GPU) on the REFERENCE's code path, for the ratio the driver computes:

* sparse side — the reference's own open GPU parameter server, HeterPS, compiled UNMODIFIED from
  /root/reference (``baseline/build_ref.py`` -> ``baseline/_ref/libref_heterps.so``) and driven the
  way ``PSGPUWrapper::PullSparse`` / ``PushSparseGrad`` drive it (``baseline/harness/ref_harness.cu``):
  cub sort + run-length merge, per-peer ``cudaMemcpyAsync`` over P2P, cuDF hash table in managed
  memory, per-key optimizer kernel, host stream syncs in between;
* dense side — what SURVEY.md §6 names as the dense half of the baseline: fp32 towers on cuBLAS
  through PyTorch, ``ncclAllReduce`` of the dense gradients from every worker thread, a separate
  scale, Adam;
* process model — the reference's: ONE process drives all N GPUs with one worker thread per GPU
  (boxps_trainer.cc:58-79).  Under torchrun rank 0 is that process and the other ranks exit.

Nothing of the product is on this path: ``paddlebox_b200`` is not imported, ``libpbx_*.so`` is not
loaded (asserted at the end).  Two things favour the reference and are deliberate: the fixed-shape
dense part of every step replays from a CUDA graph (Python's GIL would otherwise serialise eight
worker threads, which the reference's C++ workers do not suffer), and batches are packed slot-major
on the host before the timed region (the reference packs them in its feed threads).
"""
from __future__ import annotations

import ctypes as C
import json
import os
import sys
import threading
import time
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
LIB = HERE / "_ref" / "libref_heterps.so"


def unavailable(why: str) -> int:
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


# ---------------------------------------------------------------------------- synthetic data
def _mix64(k):
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * torch.tensor(-49064778989728563, dtype=torch.int64)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    k = k * torch.tensor(-4265267296055464877, dtype=torch.int64)
    k = k ^ ((k >> 33) & 0x7FFFFFFF)
    return k


def synthetic_key(i: torch.Tensor, key_seed: int = 1) -> torch.Tensor:
    x = i.to(torch.int64) * torch.tensor(-7046029254386353131, dtype=torch.int64) + key_seed
    return (_mix64(x) >> 1) & 0x7FFFFFFFFFFFFFFF


def make_host_batches(args, rank, world, n_batches, seed=1234):
    """The same generator, seeds and distributions as bench.py's product arm (so worker d sees the
    batches rank d sees there), then packed SLOT-major, the layout the reference's feed produces:
    keys [n] (slot s, instance b contiguous), lod [S*B+1] int64, dense [B, 1 + dense]."""
    g = torch.Generator().manual_seed(seed + rank)
    B, S = args.batch, args.slots
    n_total = args.keys_per_gpu * world
    out = []
    for _ in range(n_batches):
        lens = torch.randint(1, args.max_bag + 1, (B, S), generator=g)
        n = int(lens.sum())
        u = torch.rand(n, generator=g, dtype=torch.float64)
        idx = (u * u * n_total).long().clamp_(max=n_total - 1)
        keys = synthetic_key(idx, key_seed=1)                       # instance-major
        dense = torch.rand(B, 1 + args.dense, generator=g)
        dense[:, :1] = (dense[:, :1] < 0.2).float()
        # instance-major -> slot-major
        flat_lens = lens.reshape(-1)                                # [B*S], (b, s) order
        starts = torch.zeros(B * S, dtype=torch.int64)
        starts[1:] = torch.cumsum(flat_lens, 0)[:-1]
        order = torch.arange(B * S).view(B, S).t().reshape(-1)      # (s, b) order of (b, s) bags
        lens_sm = flat_lens[order]
        lod = torch.zeros(S * B + 1, dtype=torch.int64)
        lod[1:] = torch.cumsum(lens_sm, 0)
        src = torch.repeat_interleave(starts[order], lens_sm) + \
            (torch.arange(n) - torch.repeat_interleave(lod[:-1], lens_sm))
        keys_sm = keys[src]
        slot_lengths = lens.sum(0).tolist()                          # host LoD metadata per slot
        out.append((keys_sm.pin_memory(), lod.pin_memory(), dense.pin_memory(), n, slot_lengths))
    return out


# ---------------------------------------------------------------------------- one GPU worker
class Worker:
    def __init__(self, dev_index, args, lib, handle, ndev, hidden):
        self.d, self.args, self.lib, self.h, self.ndev = dev_index, args, lib, handle, ndev
        self.dev = torch.device("cuda", dev_index)
        torch.cuda.set_device(self.dev)
        self.stream = torch.cuda.Stream(device=self.dev)
        lib.ref_set_compute_stream(dev_index, C.c_void_p(self.stream.cuda_stream))
        B, S, D = args.batch, args.slots, args.dim
        self.B, self.S, self.E = B, S, 3 + D
        E = self.E
        self.in_w = S * E + args.dense
        widths = [self.in_w, *hidden, 1]
        # flat fp32 parameter / gradient / Adam-moment buffers (the reference coalesces dense
        # params into one buffer for the allreduce as well, boxps_worker.cc:373-392)
        n_par = sum(widths[i] * widths[i + 1] + widths[i + 1] for i in range(len(widths) - 1)) + 1
        g = torch.Generator().manual_seed(0)
        self.flat_p = torch.zeros(n_par, device=self.dev)
        self.flat_g = torch.zeros(n_par, device=self.dev)
        self.m = torch.zeros(n_par, device=self.dev)
        self.v = torch.zeros(n_par, device=self.dev)
        self.step_t = torch.zeros((), device=self.dev)
        self.params = []
        off = 0
        for i in range(len(widths) - 1):
            fi, fo = widths[i], widths[i + 1]
            w = self.flat_p[off:off + fi * fo].view(fo, fi)
            w.copy_((torch.rand(fo, fi, generator=g) * 2 - 1) * (1.0 / fi ** 0.5))
            wg = self.flat_g[off:off + fi * fo].view(fo, fi)
            off += fi * fo
            b = self.flat_p[off:off + fo]
            bg = self.flat_g[off:off + fo]
            off += fo
            self.params.append((w, wg, b, bg))
        self.bias, self.bias_g = self.flat_p[off:off + 1], self.flat_g[off:off + 1]
        self.leaves = []
        for w, wg, b, bg in self.params:
            for t, tg in ((w, wg), (b, bg)):
                t.requires_grad_(True)
                t.grad = tg
                self.leaves.append(t)
        self.bias.requires_grad_(True)
        self.bias.grad = self.bias_g
        # static buffers of the graphed dense part
        self.pooled = torch.zeros(S * B, E, device=self.dev, requires_grad=True)
        self.pooled.grad = torch.zeros(S * B, E, device=self.dev)
        self.feats = torch.zeros(B, args.dense, device=self.dev)
        self.label = torch.zeros(B, device=self.dev)
        self.loss = torch.zeros((), device=self.dev)
        self.graph_fb = None
        self.graph_opt = None
        self.nccl_barrier = None            # threading.Barrier shared by the workers (set by run())
        self.flush_buf = torch.empty(160 << 20, dtype=torch.uint8, device=self.dev)    # > 126 MB L2, same as the product arm

    # -- dense part: CVM transform + DeepFM (FM + deep tower) + log loss + backward ------------
    def _dense_fwd_bwd(self):
        S, B, E, D = self.S, self.B, self.E, self.args.dim
        self.flat_g.zero_()
        self.pooled.grad.zero_()
        p = self.pooled
        show_l = torch.log(p[:, 0:1] + 1.0)
        cvm = torch.cat([show_l, torch.log(p[:, 1:2] + 1.0) - show_l, p[:, 2:]], 1)     # cvm op (cvm_op.cu:29-48)
        x3 = cvm.view(S, B, E).permute(1, 0, 2)                                        # [B, S, E]
        x = torch.cat([x3.reshape(B, S * E), self.feats], 1)
        h = x
        for i, (w, _, b, _) in enumerate(self.params):
            h = torch.nn.functional.linear(h, w, b)
            if i + 1 < len(self.params):
                h = torch.relu(h)
        emb = x3[:, :, 3:]
        fm = x3[:, :, 2].sum(1) + 0.5 * ((emb.sum(1) ** 2).sum(1) - (emb ** 2).sum((1, 2)))
        logit = h[:, 0] + fm + self.bias
        loss = torch.nn.functional.binary_cross_entropy_with_logits(logit, self.label)
        loss.backward()
        self.loss.copy_(loss.detach())

    def _dense_opt(self):
        # separate scale (collective.py:240-256), then Adam (lr 1e-3, betas 0.9/0.999, eps 1e-8)
        self.flat_g.mul_(1.0 / self.ndev)
        self.step_t.add_(1.0)
        self.m.mul_(0.9).add_(self.flat_g, alpha=0.1)
        self.v.mul_(0.999).addcmul_(self.flat_g, self.flat_g, value=0.001)
        bc1 = 1.0 - 0.9 ** self.step_t
        bc2 = 1.0 - 0.999 ** self.step_t
        denom = (self.v / bc2).sqrt_().add_(1e-8)
        with torch.no_grad():
            self.flat_p.detach().addcdiv_(self.m / bc1, denom, value=-1e-3)

    def capture(self):
        torch.cuda.set_device(self.dev)
        s = torch.cuda.Stream(device=self.dev)
        s.wait_stream(self.stream)
        with torch.cuda.stream(s):
            for _ in range(3):
                self._dense_fwd_bwd()
                self._dense_opt()
        self.stream.wait_stream(s)
        torch.cuda.synchronize(self.dev)
        self.graph_fb = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_fb, stream=s, capture_error_mode="thread_local"):
            self._dense_fwd_bwd()
        self.graph_opt = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph_opt, stream=s, capture_error_mode="thread_local"):
            self._dense_opt()
        torch.cuda.synchronize(self.dev)
        # restore the state the warm-up iterations changed
        self.m.zero_(); self.v.zero_(); self.step_t.zero_()

    # -- one training step on a device-resident batch ------------------------------------------
    def step(self, batch):
        keys, lod, dense, n, slot_lengths = batch
        S, B, E = self.S, self.B, self.E
        lib, h, d = self.lib, self.h, self.d
        vals = torch.empty(n, E, device=self.dev)
        kp, vp, gp = (C.c_void_p * S)(), (C.c_void_p * S)(), (C.c_void_p * S)()
        sl = (C.c_int64 * S)(*slot_lengths)
        off = 0
        kbase, vbase = keys.data_ptr(), vals.data_ptr()
        offs = []
        for s in range(S):
            kp[s] = kbase + off * 8
            vp[s] = vbase + off * E * 4
            offs.append(off)
            off += slot_lengths[s]
        if lib.ref_pull_sparse(h, d, kp, vp, sl, S, E) != 0:
            raise RuntimeError(lib.ref_last_error().decode())
        # fused_seqpool_cvm forward, pooling half: sum-pool every bag (fused_seqpool_cvm_op.cu:35-57);
        # the CVM half is inside the graphed dense part
        self.pooled.detach().copy_(torch.segment_reduce(vals, "sum", offsets=lod, axis=0))
        self.feats.copy_(dense[:, 1:])
        self.label.copy_(dense[:, 0])
        self.graph_fb.replay()
        # fused_seqpool_cvm backward (fused_seqpool_cvm_op.cu:814-835): every key of a bag gets the
        # bag's output gradient, its first two columns carry show = 1 and clk = label
        lens = lod[1:] - lod[:-1]
        key2bag = torch.repeat_interleave(torch.arange(S * B, device=self.dev), lens, output_size=n)
        grads = self.pooled.grad.index_select(0, key2bag)
        grads[:, 0] = 1.0
        grads[:, 1] = self.label.index_select(0, key2bag % B)
        gbase = grads.data_ptr()
        for s in range(S):
            gp[s] = gbase + offs[s] * E * 4
        if lib.ref_push_sparse(h, d, gp, sl, S, E, B) != 0:
            raise RuntimeError(lib.ref_last_error().decode())
        if self.ndev > 1:
            # All workers meet before and after the allreduce (the reference has the same device for its own
            # runs, FLAGS_enable_dense_nccl_barrier).  Without it a worker that is still inside HeterComm can
            # reach a cudaMalloc (allocator miss on a new batch length) on a device whose NCCL kernel is
            # already spinning for that very worker's contribution — the classic implicit-sync deadlock,
            # which is what an 8-GPU run of this arm ran into.
            self.nccl_barrier.wait(timeout=120)
            if lib.ref_allreduce_sum(h, d, C.c_void_p(self.flat_g.data_ptr()), C.c_int64(self.flat_g.numel()),
                                     C.c_void_p(self.stream.cuda_stream)) != 0:
                raise RuntimeError(lib.ref_last_error().decode())
            self.stream.synchronize()
            self.nccl_barrier.wait(timeout=120)
        self.graph_opt.replay()
        return self.loss

    def to_device(self, hb):
        keys_h, lod_h, dense_h, n, slot_lengths = hb
        return (keys_h.to(self.dev, non_blocking=True), lod_h.to(self.dev, non_blocking=True),
                dense_h.to(self.dev, non_blocking=True), n, slot_lengths)


# ---------------------------------------------------------------------------- driver
def run(args) -> int:
    rank = int(os.environ.get("RANK", "0"))
    ndev = args.gpus
    if rank != 0:
        return 0            # the reference's process model: rank 0 drives every GPU
    if not LIB.exists():
        try:
            sys.path.insert(0, str(HERE))
            import build_ref
            build_ref.build(verbose=False)
        except Exception as e:  # noqa: BLE001
            return unavailable(f"reference HeterPS library could not be built: {str(e)[:200]}")
    if not torch.cuda.is_available() or torch.cuda.device_count() < ndev:
        return unavailable(f"needs {ndev} visible GPUs, found {torch.cuda.device_count()}")
    lib = C.CDLL(str(LIB))
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ref_last_error.restype = C.c_char_p
    lib.ref_build_synthetic.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
    lib.ref_pull_sparse.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.ref_push_sparse.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.ref_allreduce_sum.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
    lib.ref_nccl_init.argtypes = [C.c_void_p]
    lib.ref_set_compute_stream.argtypes = [C.c_int, C.c_void_p]
    lib.ref_destroy.argtypes = [C.c_void_p]

    # watchdog (armed before the table build and ncclCommInitAll, which can stall too): a run that does not finish in
    # time reports itself unavailable instead of hanging the caller
    deadline_s = float(os.environ.get("PBX_REF_DEADLINE_S", 180 + 20 * ndev))
    finished = threading.Event()

    def watchdog():
        if not finished.wait(deadline_s):
            print(json.dumps({"impl": "reference", "n_gpus": ndev,
                              "unavailable": f"reference arm did not finish within {int(deadline_s)} s at {ndev} GPUs "
                                             "(HeterComm worker threads stalled)"}), flush=True)
            os._exit(0)
    threading.Thread(target=watchdog, daemon=True).start()

    B, S = args.batch, args.slots
    hidden = tuple(int(x) for x in args.hidden.split(","))
    for d in range(ndev):
        torch.cuda.set_device(d)
        torch.zeros(1, device=f"cuda:{d}")
    torch.cuda.set_device(0)
    h = lib.ref_create(ndev, args.dim, S, 1)
    t0 = time.time()
    counts = (C.c_int64 * ndev)()
    if lib.ref_build_synthetic(h, args.keys_per_gpu * ndev, 1, 1, counts) != 0:
        return unavailable("reference table build failed: " + lib.ref_last_error().decode()[:200])
    fill_s = time.time() - t0
    if ndev > 1 and lib.ref_nccl_init(h) != 0:
        return unavailable("ncclCommInitAll failed: " + lib.ref_last_error().decode()[:200])

    workers = [Worker(d, args, lib, h, ndev, hidden) for d in range(ndev)]
    nccl_barrier = threading.Barrier(ndev)
    for w in workers:
        w.nccl_barrier = nccl_barrier
        w.capture()
    host_batches = [make_host_batches(args, d, ndev, args.n_batches) for d in range(ndev)]
    barrier = threading.Barrier(ndev)
    results = [None] * ndev
    errors = []
    clock_lines = []

    def clocks_thread(stop):
        import subprocess
        q = ("index,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200",
                                  "-i", "0"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:  # noqa: BLE001
            return
        import selectors
        sel = selectors.DefaultSelector()
        sel.register(p.stdout, selectors.EVENT_READ)
        while not stop.is_set():
            if sel.select(timeout=0.1):
                ln = p.stdout.readline()
                if ln:
                    clock_lines.append(ln.strip())
        p.terminate()

    def work(d):
        try:
            w = workers[d]
            torch.cuda.set_device(w.dev)
            torch.cuda.set_stream(w.stream)
            hbs = host_batches[d]
            dev_batches = [w.to_device(hb) for hb in hbs]
            torch.cuda.synchronize(w.dev)
            for i in range(args.warmup):
                w.step(dev_batches[i % len(dev_batches)])
            torch.cuda.synchronize(w.dev)
            barrier.wait()
            # ---- device-timed steps, inputs resident, L2 flushed between steps
            evs = []
            for i in range(args.steps):
                w.flush_buf.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                loss = w.step(dev_batches[(args.warmup + i) % len(dev_batches)])
                b.record()
                evs.append((a, b))
            torch.cuda.synchronize(w.dev)
            barrier.wait()
            dev_ms = sum(a.elapsed_time(b) for a, b in evs)
            loss_val = float(loss)
            # ---- end to end: every step copies its inputs from pinned host memory and reads the loss back
            h2d = d2h = 0
            for i in range(2):
                w.step(w.to_device(hbs[i % len(hbs)]))
            torch.cuda.synchronize(w.dev)
            barrier.wait()
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record()
            for i in range(args.steps):
                hb = hbs[(2 + i) % len(hbs)]
                db = w.to_device(hb)
                h2d += sum(t.numel() * t.element_size() for t in hb[:3])
                lv = w.step(db).item()          # device -> host read of the step's loss
                d2h += 4
                assert lv == lv
                w.flush_buf.zero_()
            eb.record()
            torch.cuda.synchronize(w.dev)
            barrier.wait()
            results[d] = (dev_ms, ea.elapsed_time(eb), loss_val, h2d // args.steps, d2h // args.steps)
        except Exception as e:  # noqa: BLE001
            # the other workers are (or will be) waiting for this one inside NCCL or at a barrier: there is no
            # clean way to unwind them, so report and leave
            import traceback
            print(f"worker {d}: {e}\n{traceback.format_exc()}", file=sys.stderr, flush=True)
            print(json.dumps({"impl": "reference", "n_gpus": ndev,
                              "unavailable": f"reference arm failed at run time on worker {d}: {str(e)[:160]}"}),
                  flush=True)
            os._exit(0)

    stop = threading.Event()
    ct = threading.Thread(target=clocks_thread, args=(stop,), daemon=True)
    ct.start()
    wall0 = time.time()
    threads = [threading.Thread(target=work, args=(d,)) for d in range(ndev)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    wall = time.time() - wall0
    finished.set()
    stop.set()
    ct.join(timeout=2)
    if errors:
        print(errors[0], file=sys.stderr)
        return unavailable("reference arm failed at run time: " + errors[0].splitlines()[0][:200])

    with open("/proc/self/maps") as f:
        maps = f.read()
    assert "libpbx_" not in maps and "paddlebox_b200" not in sys.modules, "product code on the reference path"

    dev_ms = max(r[0] for r in results)
    e2e_ms = max(r[1] for r in results)
    sm, mx, reasons = [], [], set()
    for ln in clock_lines:
        f = [x.strip() for x in ln.split(",")]
        if len(f) >= 7:
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nme, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
    sm.sort()
    gb = B * ndev
    out = {
        "metric": "instances/sec (DeepFM sparse-CTR training, device-timed, max over ranks)",
        "value": round(gb * args.steps / (dev_ms / 1e3), 1), "unit": "instances/s", "n_gpus": ndev,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dev_ms / args.steps, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "impl": "reference",
        "config": {"model": "DeepFM", "global_batch": gb, "seq_len": None,
                   "parallelism": f"dp{ndev}+sparse-shard{ndev} (one process, one worker thread per GPU)",
                   "batch_per_gpu": B, "sparse_slots": S, "dense_features": args.dense, "emb_dim": args.dim,
                   "bag_len": f"U[1,{args.max_bag}]",
                   "avg_keys_per_batch_per_gpu": int(sum(hb[3] for hb in host_batches[0]) / len(host_batches[0])),
                   "table_keys_per_gpu": args.keys_per_gpu, "table_keys_total": args.keys_per_gpu * ndev,
                   "table_keys_on_gpu": [int(c) for c in counts],
                   "mlp": [workers[0].in_w, *hidden, 1], "sparse_optimizer": "adagrad", "dense_optimizer": "adam",
                   "table_dtype": "fp32", "tower_dtype": "fp32 (the reference's towers are fp32 cuBLAS)",
                   "sparse_engine": "reference HeterPS (heter_ps/*, unmodified, sm_100 SIMT build), "
                                    "FLAGS_gpugraph_dedup_pull_push_mode=%s" %
                                    os.environ.get("FLAGS_gpugraph_dedup_pull_push_mode", "0 (default)"),
                   "hash_load_factor": 0.75, "l2": "flushed between timed steps (160 MB write)",
                   "table_fill_s": round(fill_s, 2), "cuda_graph": "dense fwd/bwd and optimizer only"},
        "e2e": {"value": round(gb * args.steps / (e2e_ms / 1e3), 1), "unit": "instances/s",
                "ms_per_step": round(e2e_ms / args.steps, 4), "h2d_bytes_per_step": int(results[0][3]),
                "d2h_bytes_per_step": int(results[0][4]),
                "timed_region": "one CUDA-event pair per worker around all K steps incl. H2D of every batch from "
                                "pinned memory, the step, the per-step loss read and the per-step L2 flush"},
        "gpu_launches": None,
        "clocks": {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                   "reasons": sorted(reasons), "samples": len(sm)},
        "loss": results[0][2], "wall_s": round(wall, 2),
    }
    print(json.dumps(out))
    sys.stdout.flush()
    # skip the destructors of the reference objects (the managed-memory tables take long to free)
    os._exit(0)
