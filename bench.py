#!/usr/bin/env python3
"""bench.py -- SGD row*iterations/s of the per-micro-batch learning path on N B200s.

A "step" is one micro-batch: the hot path (numIterations fused SGD-iteration passes + their
fixed-order reduce/update) over a synthetic dense micro-batch already resident in HBM.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference ...                      (the CPU restatement, all host cores)

Workloads (BASELINE.json configs):
  dense_1m    2^20 rows/GPU x 1004 features, 20 iterations  (configs[1]; the default)
  dense_16m   2^24 rows/GPU x 1004 features, 50 iterations  (configs[2], HBM-roofline config; no e2e:
              134.75 GB does not stage through host memory)
  dense_d128  2^24 rows/GPU x 128 features, 50 iterations (the narrow-row shape of configs[3])
  tweets_1m   10^6 synthetic tweets/GPU (SURVEY 8d config 1 recipe) -> device featuriser (HashingTF(1000) + 4) ->
              CSR of the rows that pass the retweet filter (~14 %), 50 iterations, reference step size 0.005
  csr_2m      2^21 rows/GPU, 2^20+4 columns, ~24 nnz/row CSR, 50 iterations (configs[4] per GPU)
Multi-GPU: every rank holds the same number of rows (weak scaling); per iteration the ranks exchange
their (gradSum, lossSum, count) shard vectors with one NCCL all-gather.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "dense_1m": dict(rows=1 << 20, d=1004, iters=20, step=0.1),
    "dense_16m": dict(rows=1 << 24, d=1004, iters=50, step=0.1),
    # configs[3] strong-scaling shape (2^27 x 128 over the GPUs): here 2^24 narrow rows per GPU
    "dense_d128": dict(rows=1 << 24, d=128, iters=50, step=0.1),
    # configs[4] per GPU: 2^21 rows (2^24 over 8 GPUs), 2^20 hashed + 4 numeric columns, ~24 nnz/row, uniform columns
    "csr_2m": dict(rows=1 << 21, d=(1 << 20) + 4, d_text=1 << 20, iters=50, step=0.005, csr=True),
}
WORKLOADS["tweets_1m"] = dict(rows=1000000, d=1004, iters=50, step=0.005, csr=True, tweets=True, buckets=1000)
# configs[4], distribution A: the same tweets hashed into 2^20 buckets -- char bigrams of mostly-ASCII text land in a few
# thousand of them, i.e. extreme column skew
WORKLOADS["tweets_1m_wide"] = dict(rows=1000000, d=(1 << 20) + 4, iters=50, step=0.005, csr=True, tweets=True, buckets=1 << 20)
METRIC = "sgd_row_iterations_per_s"
UNIT = "row*iter/s"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def synth_tweets(n, seed):
    """Vectorised form of the SURVEY.md 8d config-1 recipe: lower-case text of U[2,140] tokens over letters, digits,
    punctuation, 16 Latin-1 accented letters and 8 emoji (two UTF-16 units each); retweet ~ Bernoulli(0.7); log-uniform
    counts; created within 7 days before a fixed now.  -> (TweetBatch columns, now_ms)"""
    import numpy as np
    rng = np.random.default_rng(seed)
    now = 1464739200000
    one = [ord(c) for c in "abcdefghijklmnopqrstuvwxyz0123456789 #@.,!?:/" + "\u00e0\u00e1\u00e2\u00e3\u00e4\u00e5\u00e7\u00e8\u00e9\u00ea\u00eb\u00ec\u00ed\u00f1\u00f6\u00fc"]
    emoji = [0x1F600, 0x1F601, 0x1F602, 0x1F923, 0x1F60D, 0x1F44D, 0x1F525, 0x1F389]
    first = np.array(one + [0xD800 + ((c - 0x10000) >> 10) for c in emoji], np.uint16)
    second = np.array([0] * len(one) + [0xDC00 + ((c - 0x10000) & 0x3FF) for c in emoji], np.uint16)
    tlen = np.array([1] * len(one) + [2] * len(emoji), np.int64)
    ntok = rng.integers(2, 141, n)
    tok = rng.integers(0, len(first), int(ntok.sum()))
    ulen = tlen[tok]
    pos = np.cumsum(ulen) - ulen
    units = np.zeros(int(ulen.sum()), np.uint16)
    units[pos] = first[tok]
    two = ulen == 2
    units[pos[two] + 1] = second[tok[two]]
    tok_end = np.cumsum(ntok)
    cum_units = np.concatenate([[0], np.cumsum(ulen)])
    text_off = np.concatenate([[0], cum_units[tok_end]]).astype(np.int64)
    lu = lambda lo, hi: np.floor(np.exp(rng.uniform(np.log(lo), np.log(hi), n))).astype(np.int64)
    cols = dict(text_utf16=units, text_off=text_off, retweet_count=lu(1, 1e5), followers=lu(1, 1e8), favourites=lu(1, 1e5) - 1,
                friends=lu(1, 1e5) - 1, created_ms=now - rng.integers(0, 7 * 86400_000, n), is_retweet=(rng.random(n) < 0.7))
    return cols, now


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_leg(wl, steps, warmup, budget_s=15.0, threads=None):
    """The CPU restatement (oracle/twtml_oracle.c, OpenMP over row partitions like Spark local[N]) on a
    bounded sample of the workload: same width, same iteration count, fewer rows (dense: 2^17 rows;
    CSR: 2^18 rows; tweets: 10^5 tweets, featurised by the restatement inside the timed step)."""
    from oracle import oracle_c
    import numpy as np
    oracle_c.build()
    cores = threads or oracle_c.max_threads()
    d, iters = wl["d"], wl["iters"]
    w0 = np.zeros(d)
    common = dict(step_size=wl["step"], num_iterations=iters, convergence_tol=0.0, partitions=cores, threads=cores)
    if wl.get("tweets"):
        n_t = 100000
        cols, now = synth_tweets(n_t, 20160601)
        c = {k: np.ascontiguousarray(v) for k, v in cols.items()}

        def step():
            rp, ci, va, la = oracle_c.featurize(c["text_utf16"], c["text_off"], c["retweet_count"], c["followers"], c["favourites"],
                                                c["friends"], c["created_ms"], c["is_retweet"].astype(np.uint8), now, 100, 1000, wl["buckets"])
            oracle_c.sgd(la, d, w0, csr=(rp, ci, va), **common)
            return len(la)
        what = "%d of %d tweets (featuriser + %d iterations on the rows that pass the filter)" % (n_t, wl["rows"], iters)
    elif wl.get("csr"):
        rows = 1 << 18
        rp, ci, va, la = oracle_c.synth_csr(rows, wl["d_text"], seed=42)

        def step():
            oracle_c.sgd(la, d, w0, csr=(rp, ci, va), **common)
            return rows
        what = "%d of %d rows x %d columns CSR, %d iterations" % (rows, wl["rows"], d, iters)
    else:
        rows = 1 << 17
        X, y = oracle_c.synth_dense(rows, d, seed=42)

        def step():
            oracle_c.sgd(y, d, w0, X=X, **common)
            return rows
        what = "%d of %d rows x %d features, %d iterations" % (rows, wl["rows"], d, iters)
    t0 = time.perf_counter()
    step()                                              # also the first warm-up step
    once = time.perf_counter() - t0
    k = max(1, min(steps, int(budget_s / max(once, 1e-3))))
    for _ in range(min(max(0, warmup - 1), 2)):
        step()
    t0 = time.perf_counter()
    for _ in range(k):
        units = step()
    dt = time.perf_counter() - t0
    value = units * iters * k / dt
    sample = "%s, %d steps, OpenMP static row partitions on %d threads" % (what, k, cores)
    return dict(value=value, unit=UNIT, cores=cores, kind="port", sample=sample), dt / k * 1e3, k


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, ms, k = cpu_leg(wl, args.steps, args.warmup)
    cb2, _, _ = cpu_leg(wl, 2, 1, budget_s=8.0, threads=2)            # README.md:55 of the reference: --master local[2]
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": k,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "rows_per_gpu": wl["rows"], "features": wl["d"], "sgd_iterations": wl["iters"],
                       "note": "CPU restatement of the reference's MLlib 1.6.1 path (the Scala/Spark reference cannot run here: no JVM)"},
            "cpu_baseline": cb, "cpu_baseline_local2": cb2,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="dense_1m", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU (testing)")
    ap.add_argument("--iters", type=int, default=0, help="override SGD iterations per micro-batch (testing)")
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--e2e-sync", action="store_true", help="blocking ingest in the e2e leg (A/B against the pipelined default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.rows:
        wl["rows"] = args.rows
    if args.iters:
        wl["iters"] = args.iters
    if args.impl == "reference":
        return run_reference(args, wl)
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    import torch.distributed as dist

    import twtml_b200 as tw

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N > 1 with torch.distributed.run" % (args.gpus, world))
    torch.cuda.set_device(local)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ids = [tw.Context.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        nccl_id = ids[0]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n, d, iters = wl["rows"], wl["d"], wl["iters"]
    ctx = tw.Context(local, rank, world, nccl_id)
    ctx.profile(True)
    model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(iters).setStepSize(wl["step"])
             .setMiniBatchFraction(1.0).setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(d)))
    csr = bool(wl.get("csr"))
    tweets = None
    if wl.get("tweets"):
        cols, now_ms = synth_tweets(n, 20160601 + rank)
        tweets = tw.TweetBatch.from_arrays(**cols)
        batch = ctx.batch_from_tweets(tweets, now_ms, 100, 1000, wl["buckets"])   # filter + featurise on the GPU
        n_tweets, n = n, batch.n                                     # the metric counts the rows that train
    elif csr:
        batch = ctx.batch_synth_csr(n, wl["d_text"], seed=42, row0=rank * n)
    else:
        batch = ctx.batch_synth_dense(n, d, seed=42, row0=rank * n)  # resident in HBM before the timed region

    for _ in range(args.warmup):
        model.trainOnBatch(batch)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms = kern_ms = 0.0
    launches = kern_launches = 0
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        model.trainOnBatch(batch)
        p = ctx.last_profile()
        dev_ms += p.loop_ms
        kern_ms += p.iter_kernel_ms
        kern_launches += p.iter_launches
        launches += p.iter_launches + p.other_launches
        nccl_calls = p.nccl_calls
        kbytes = p.iter_kernel_bytes
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = max_over_ranks(dev_ms)
    wall_ms = max_over_ranks(wall_ms)
    kern_ms_per_launch = max_over_ranks(kern_ms / max(kern_launches, 1))
    assert len(model.lastLossHistory) == iters, "iterations were skipped"
    value = world * n * iters * args.steps / (dev_ms * 1e-3)

    # ---- end to end through the C ABI with HOST buffers: ingest (H2D) + stats + train + results (D2H)
    e2e = None
    if (args.workload != "dense_16m" or args.rows) and args.e2e_steps > 0:
        def pin(a):  # page-locked copy of a host array (the library then DMAs straight from it)
            raw = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory().numpy()
            return raw.view(a.dtype)

        if tweets is not None:
            h2d = sum(getattr(tweets, k).nbytes for k in ("text_utf16", "text_off", "retweet_count", "followers", "favourites",
                                                           "friends", "created_ms", "is_retweet"))
            ptweets = tw.TweetBatch.from_arrays(*(pin(getattr(tweets, k)) for k in
                                                 ("text_utf16", "text_off", "retweet_count", "followers", "favourites", "friends",
                                                  "created_ms", "is_retweet")))
            make = lambda: ctx.batch_from_tweets(ptweets, now_ms, 100, 1000, wl["buckets"])
        elif csr:
            host = tuple(pin(a) for a in batch.export_csr())           # page-locked host copy of this rank's rows
            h2d = sum(a.nbytes for a in host)
            make = lambda: ctx.batch_from_csr(n, d, *host)
        else:
            Xh, yh = batch.export_dense()
            Xp = torch.from_numpy(Xh).pin_memory().numpy()
            del Xh
            yp = torch.from_numpy(yh).pin_memory().numpy()
            h2d = n * (d + 1) * 8
            # page-locked buffers: the library DMAs straight from them on its copy stream, so the ingest of
            # step t+1 is issued before step t computes and overlaps it (streaming ingest)
            make = (lambda: ctx.batch_from_dense(Xp, yp)) if args.e2e_sync else (lambda: ctx.batch_from_dense_async(Xp, yp))
        model.setInitialWeights(tw.Vectors.zeros(d))
        t_ingest = t_step = 0.0

        def run(k):
            nonlocal t_ingest, t_step
            nxt = make()
            for i in range(k):
                b = nxt
                t0 = time.perf_counter()
                nxt = make() if i + 1 < k else None        # every step's rows cross PCIe inside the timed region
                t1 = time.perf_counter()
                st, _, _ = model.step(b, want_arrays=False)
                w = model.latestModel().weights.toArray()
                t_ingest += t1 - t0
                t_step += time.perf_counter() - t1
                b.free()
            return st, w

        run(1)
        t_ingest = t_step = 0.0
        barrier()
        t0 = time.perf_counter()
        st, w = run(args.e2e_steps)
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": world * n * iters * args.e2e_steps / e2e_s, "unit": UNIT,
               "h2d_bytes_per_step": int(world * h2d),
               "d2h_bytes_per_step": int(world * (d * 8 + iters * 8 + 80)),
               "steps": args.e2e_steps, "ingest_ms_per_step": t_ingest / args.e2e_steps * 1e3,
               "compute_ms_per_step": t_step / args.e2e_steps * 1e3,
               "h2d_GBps_effective": h2d * args.e2e_steps / e2e_s / 1e9,
               "includes": "host buffers -> HBM ingest (H2D%s), stats + %d iterations, weights/loss/stats back"
                           % (" + transpose build" if csr else ", pinned, next step's copy overlapping this step's compute", iters)}

    narrow = bool(csr) and d <= 2048 and not os.environ.get("TWTML_NO_NARROW")  # csr_narrow.cuh: kNarrowMaxD
    if rank == 0:
        peak, peak_src = measured_peak()
        achieved = kbytes / (kern_ms_per_launch * 1e-3) / 1e9
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "iter_kernel_traffic.json")) as f:
                t = json.load(f)
                for ent in (t if isinstance(t, list) else [t]):
                    if ent.get("rows") == n and ent.get("features") == d and bool(ent.get("csr", False)) == bool(csr):
                        traffic = ent["dram_bytes_per_launch"]
        except Exception:
            pass
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": args.workload, "rows_per_gpu": n, "features": d, "sgd_iterations": iters,
                           "convergence_tol": 0.0, **({"tweets_per_gpu": n_tweets, "nnz_per_gpu": batch.nnz} if tweets is not None else {}), "parallelism": "rows sharded x%d, one exchange of the shard partials per iteration" % world,
                           "l2": ("inputs (%.3f GB/GPU) about the size of L2, fetched with an evict-first policy: ncu shows every launch reading them from HBM (profiles/); no flush"
                                  % ((batch.nnz * 12 + n * 16) / 1e9)) if narrow else
                                 "inputs (%.2f GB/GPU) larger than L2; no flush needed" % ((batch.nnz * 12 * 2 + n * 24) / 1e9 if csr else n * (d + 2) * 8 / 1e9),
                           "timing": "CUDA events on the library stream around each step's iteration loop, max over ranks"},
                "wall_ms_per_step": wall_ms / args.steps,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "peak_source": peak_src,
                             "kernel": ("k_csr_narrow<TRAIN> (single pass over the CSR)" if narrow else "k_csr_rows + k_csc_windows + k_csc_edges (one read of the CSR credited)") if csr else ("k_dense_iter<1,TRAIN,RL1>" if d + 1 > 992 else "k_dense_iter<1,TRAIN,row lanes>"),
                             "kernel_ms": kern_ms_per_launch, "bytes_per_launch": int(kbytes),
                             "frac_of_nominal_8TBs": achieved / 8000.0},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
                "exchange": ("none (1 GPU)" if world == 1 else (("NCCL all-to-all of gradient slices + all-gather of w, %d calls per step" if nccl_calls == 2 * iters else
                                                                  "NCCL all-gather, %d calls per step") % nccl_calls if nccl_calls else
                                                                "peer-memory stores + flags inside the reduce/update kernels (0 NCCL calls per step)"))}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _, _ = cpu_leg(wl, 3, 1)
        print(json.dumps(line), flush=True)
    batch.free()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
