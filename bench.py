#!/usr/bin/env python3
"""bench.py -- SGD row*iterations/s of the per-micro-batch learning path on N B200s.

A "step" is one micro-batch: the hot path (numIterations fused SGD-iteration passes + their
fixed-order reduce/update) over a synthetic micro-batch already resident in HBM.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference ...                      (the CPU restatement, all host cores)

Workloads (BASELINE.json configs):
  dense_16m   2^24 rows/GPU x 1004 features, 50 iterations  (configs[2], the HBM-roofline config the metric is quoted
              on: THE DEFAULT).  Its end-to-end leg streams the 135 GB batch from one 2^20-row page-locked host buffer,
              re-sent 16 times per step through twtml_batch_dense_begin / _append (every byte crosses the host link
              inside the timed region; the host never holds more than 8.4 GB).
  dense_1m    2^20 rows/GPU x 1004 features, 20 iterations  (configs[1])
  dense_d128  2^24 rows/GPU x 128 features, 50 iterations   (the narrow-row shape of configs[3])
  tweets_1m   10^6 synthetic tweets/GPU (SURVEY 8d config 1 recipe) -> device featuriser (HashingTF(1000) + 4) ->
              CSR of the rows that pass the retweet filter (~14 %), 50 iterations, reference step size 0.005  (configs[0])
  csr_2m      2^21 rows/GPU, 2^20+4 columns, ~24 nnz/row CSR, 50 iterations (configs[4] per GPU)
The JSON line is the default workload's; at N = 1 it also carries "sub_records" with the same measurements
(value, roofline, e2e) for dense_1m, tweets_1m and csr_2m (--sub none turns them off).
Multi-GPU: every rank holds the same number of rows (weak scaling); per iteration the ranks exchange
their (gradSum, lossSum, count) shard vectors (peer-memory stores, NCCL as the fallback).
The "e2e" object also carries "step_ms", this rank's wall clock of every end-to-end step (a stall shows there, not in
the mean).  For tweets, "h2d_bytes_per_step" counts what crosses the host link -- the copied columns plus the text of the
tweets that pass the filter, which the featuriser reads in place from the page-locked column -- and
"host_input_bytes_per_step" the whole input as it lies in host memory.
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
DEFAULT_WORKLOAD = "dense_16m"
DEFAULT_SUBS = "dense_1m,tweets_1m,csr_2m"
METRIC = "sgd_row_iterations_per_s"
UNIT = "row*iter/s"
CHUNK_ROWS = 1 << 20          # rows of the page-locked host buffer the dense e2e legs stream from


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def workload_config(name, wl, world):
    """What both arms (ours and --impl reference) run: the same keys and values on both lines."""
    return {"workload": name, "rows_per_gpu": wl["rows"], "features": wl["d"], "sgd_iterations": wl["iters"],
            "step_size": wl["step"], "convergence_tol": 0.0, "mini_batch_fraction": 1.0,
            "layout": "csr" if wl.get("csr") else "dense", "parallelism": "rows sharded x%d" % world}


def synth_tweets(n, seed):
    """Vectorised form of the SURVEY.md 8d config-1 recipe: lower-case text of U[2,140] tokens over letters, digits,
    punctuation, 16 Latin-1 accented letters and 8 emoji (two UTF-16 units each); retweet ~ Bernoulli(0.7); log-uniform
    counts; created within 7 days before a fixed now.  -> (TweetBatch columns, now_ms)"""
    import numpy as np
    rng = np.random.default_rng(seed)
    now = 1464739200000
    one = [ord(c) for c in "abcdefghijklmnopqrstuvwxyz0123456789 #@.,!?:/" + "àáâãäåçèéêëìíñöü"]
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


def host_cores():
    """Cores this process may run on -- NOT OMP_NUM_THREADS, which torchrun sets to 1 for its workers."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def bind_to_gpu_numa_node(local):
    """Run this process on the CPUs of the NUMA node the GPU hangs off, so that the page-locked host buffers it
    allocates next are local to the GPU's PCIe root (first-touch placement).  Returns a description, or None."""
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(local), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bdf = out[-12:] if len(out) >= 12 else out          # 00000000:1b:00.0 -> 0000:1b:00.0
        with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
            node = int(f.read().strip())
        if node < 0:
            return {"node": node, "bound": False}
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return {"node": node, "bound": False}
        os.sched_setaffinity(0, allowed)
        return {"node": node, "bound": True, "cpus": len(allowed)}
    except Exception as e:                                     # no sysfs / no nvidia-smi: leave the process where it is
        return {"bound": False, "why": str(e)[:80]}


def cpu_leg(wl, steps, warmup, threads=None, max_seconds=150.0):
    """The CPU restatement (oracle/twtml_oracle.c, OpenMP over row partitions like Spark local[N]) on a
    bounded sample of the workload: same width, same iteration count, fewer rows (dense: 2^17 rows;
    CSR: 2^18 rows; tweets: 10^5 tweets, featurised by the restatement inside the timed step).  Runs `steps` timed
    steps after `warmup` untimed ones unless that would take longer than max_seconds (then fewer, and says so)."""
    from oracle import oracle_c
    import numpy as np
    oracle_c.build()
    cores = threads or host_cores()
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
        X, y = oracle_c.synth_dense(rows, d, seed=42, threads=cores)

        def step():
            oracle_c.sgd(y, d, w0, X=X, **common)
            return rows
        what = "%d of %d rows x %d features, %d iterations" % (rows, wl["rows"], d, iters)
    t0 = time.perf_counter()
    step()                                              # also the first warm-up step
    once = time.perf_counter() - t0
    k = max(1, min(steps, int(max_seconds / max(once, 1e-3)) - warmup))
    for _ in range(max(0, warmup - 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(k):
        units = step()
    dt = time.perf_counter() - t0
    value = units * iters * k / dt
    sample = "%s, %d steps, OpenMP static row partitions on %d threads" % (what, k, cores)
    return dict(value=value, unit=UNIT, cores=cores, kind="port", sample=sample), dt / k * 1e3, k


def run_reference(args, name, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, ms, k = cpu_leg(wl, args.steps, args.warmup)
    cb2, _, _ = cpu_leg(wl, 2, 1, threads=2, max_seconds=20.0)            # README.md:55 of the reference: --master local[2]
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": k,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_config(name, wl, args.gpus),
            "note": "CPU restatement of the reference's MLlib 1.6.1 path (the Scala/Spark reference cannot run here: no JVM), "
                    "timed on a bounded row sample of the workload (cpu_baseline.sample)",
            "cpu_baseline": cb, "cpu_baseline_local2": cb2,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def traffic_entry(n, d, csr):
    """DRAM bytes per launch of the iteration kernel from the committed ncu capture of this kernel at this shape
    (profiles/iter_kernel_traffic.json); None when no capture of this shape exists."""
    try:
        with open(os.path.join(ROOT, "profiles", "iter_kernel_traffic.json")) as f:
            t = json.load(f)
        for ent in (t if isinstance(t, list) else [t]):
            if ent.get("rows") == n and ent.get("features") == d and bool(ent.get("csr", False)) == bool(csr):
                return ent["dram_bytes_per_launch"], ent.get("source")
    except Exception:
        pass
    return None, None


class Bench:
    """One process's (one rank's) measurement state."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        import twtml_b200 as tw
        self.torch, self.dist, self.tw, self.args = torch, dist, tw, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N > 1 with torch.distributed.run" % (args.gpus, self.world))
        self.numa = None if args.no_numa_bind else bind_to_gpu_numa_node(self.local)
        torch.cuda.set_device(self.local)
        nccl_id = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            ids = [tw.Context.nccl_unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            nccl_id = ids[0]
        self.ctx = tw.Context(self.local, self.rank, self.world, nccl_id)
        self.ctx.profile(True)
        self.chunk = None          # page-locked (X, y) of CHUNK_ROWS x 1004, shared by the dense e2e legs

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def pin(self, a):
        """Page-locked copy of a host array (the library then DMAs straight from it)."""
        import numpy as np
        raw = self.torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory().numpy()
        return raw.view(a.dtype).reshape(a.shape)

    def host_chunk(self, d, rows):
        """>= rows x d page-locked rows + labels: generated on the device, copied out once, then page-locked."""
        if self.chunk is None or self.chunk[0].shape[1] != d or self.chunk[0].shape[0] < rows:
            self.chunk = None
            g = self.ctx.batch_synth_dense(rows, d, seed=42, row0=self.rank * rows)
            Xh, yh = g.export_dense()
            g.free()
            Xp = self.torch.from_numpy(Xh).pin_memory().numpy()
            del Xh
            self.chunk = (Xp, self.torch.from_numpy(yh).pin_memory().numpy())
        return self.chunk

    # ------------------------------------------------------------------------------------------------
    def run(self, name, wl, steps, warmup, e2e_steps, sample_clocks):
        import numpy as np
        tw, ctx, world, rank = self.tw, self.ctx, self.world, self.rank
        n, d, iters = wl["rows"], wl["d"], wl["iters"]
        model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setNumIterations(iters).setStepSize(wl["step"])
                 .setMiniBatchFraction(1.0).setConvergenceTol(0.0).setInitialWeights(tw.Vectors.zeros(d)))
        csr = bool(wl.get("csr"))
        tweets = n_tweets = None
        if wl.get("tweets"):
            cols, now_ms = synth_tweets(n, 20160601 + rank)
            tweets = tw.TweetBatch.from_arrays(**cols)
            batch = ctx.batch_from_tweets(tweets, now_ms, 100, 1000, wl["buckets"])   # filter + featurise on the GPU
            n_tweets, n = n, batch.n                                     # the metric counts the rows that train
        elif csr:
            batch = ctx.batch_synth_csr(n, wl["d_text"], seed=42, row0=rank * n)
        else:
            batch = ctx.batch_synth_dense(n, d, seed=42, row0=rank * n)  # resident in HBM before the timed region
        nnz = batch.nnz

        for _ in range(warmup):
            model.trainOnBatch(batch)
        sampler = ClockSampler(self.local) if (sample_clocks and rank == 0) else None
        if sampler:
            sampler.start()
        dev_ms = kern_ms = 0.0
        launches = kern_launches = 0
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            model.trainOnBatch(batch)
            p = ctx.last_profile()
            dev_ms += p.loop_ms
            kern_ms += p.iter_kernel_ms
            kern_launches += p.iter_launches
            launches += p.iter_launches + p.other_launches
            nccl_calls = p.nccl_calls
            kbytes = p.iter_kernel_bytes
        self.barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        clocks = sampler.stop() if sampler else None
        dev_ms = self.max_over_ranks(dev_ms)
        wall_ms = self.max_over_ranks(wall_ms)
        kern_ms_per_launch = self.max_over_ranks(kern_ms / max(kern_launches, 1))
        assert len(model.lastLossHistory) == iters, "iterations were skipped"
        value = world * n * iters * steps / (dev_ms * 1e-3)
        import hashlib
        w_now = np.ascontiguousarray(model.latestModel().weights.toArray())
        weights_sha = hashlib.sha256(w_now.tobytes()).hexdigest()[:16]   # equal across N when the N ranks hold the same global batch

        # ---- end to end through the C ABI with HOST buffers: ingest (H2D) + stats + train + results (D2H)
        e2e = None
        if e2e_steps > 0:
            streamed = False
            if tweets is not None:
                names = ("text_utf16", "text_off", "retweet_count", "followers", "favourites", "friends", "created_ms", "is_retweet")
                h2d = host_input = sum(getattr(tweets, k).nbytes for k in names)
                ptweets = tw.TweetBatch.from_arrays(*(self.pin(getattr(tweets, k)) for k in names))
                make = lambda: ctx.batch_from_tweets(ptweets, now_ms, 100, 1000, wl["buckets"], wait=self.args.e2e_sync)
                how = "page-locked tweet columns -> device featuriser"
                if 2 * n <= n_tweets and not os.environ.get("TWTML_NO_TEXT_IN_PLACE"):
                    # twtml_sparse.cu tweets_ingest: the text column stays in host memory; the featuriser reads the text of
                    # the tweets that pass the filter over the host link, inside the timed region
                    kept = (tweets.is_retweet != 0) & (tweets.retweet_count >= 100) & (tweets.retweet_count <= 1000)
                    assert int(kept.sum()) == n
                    kept_text = 2 * int(np.diff(tweets.text_off)[kept].sum())
                    h2d = host_input - tweets.text_utf16.nbytes + kept_text
                    how = ("page-locked tweet columns -> device featuriser; the text column (%d B) is read in place, %d B of it "
                           "(the kept tweets') cross the link" % (tweets.text_utf16.nbytes, kept_text))
            elif csr:
                host = tuple(self.pin(a) for a in batch.export_csr())           # page-locked host copy of this rank's rows
                h2d = sum(a.nbytes for a in host)
                make = lambda: ctx.batch_from_csr(n, d, *host, wait=self.args.e2e_sync)
                how = "page-locked CSR arrays" + ("" if d <= 2048 else " + transpose build")
            else:
                batch.free()                                                    # two 135 GB batches do not fit
                batch = None
                Xp, yp = self.host_chunk(d, min(n, CHUNK_ROWS))
                h2d = n * (d + 1) * 8
                if n <= CHUNK_ROWS:
                    Xv, yv = Xp[:n], yp[:n]
                    make = (lambda: ctx.batch_from_dense(Xv, yv)) if self.args.e2e_sync else (lambda: ctx.batch_from_dense_async(Xv, yv))
                    how = "page-locked rows, flat 64 MB copies + k_repitch"
                else:
                    streamed = True

                    def make():
                        b = ctx.batch_dense_begin(n, d)
                        for r0 in range(0, n, CHUNK_ROWS):
                            nr = min(CHUNK_ROWS, n - r0)
                            b.append(Xp[:nr], yp[:nr])
                        return b
                    how = ("one %d-row page-locked buffer re-sent %d times per step through twtml_batch_dense_begin/_append"
                           % (CHUNK_ROWS, (n + CHUNK_ROWS - 1) // CHUNK_ROWS))
            model.setInitialWeights(tw.Vectors.zeros(d))
            t_ingest = t_step = 0.0
            step_ms = []

            def run(k):
                nonlocal t_ingest, t_step
                nxt = make()
                for i in range(k):
                    b = nxt
                    t0 = time.perf_counter()
                    # every step's rows cross the host link inside the timed region; where HBM holds two batches the next
                    # step's ingest is issued before this step computes and overlaps it
                    nxt = make() if (i + 1 < k and not streamed) else None
                    t1 = time.perf_counter()
                    st, _, _ = model.step(b, want_arrays=False)
                    w = model.latestModel().weights.toArray()
                    t_ingest += t1 - t0
                    t_step += time.perf_counter() - t1
                    step_ms.append((time.perf_counter() - t0) * 1e3)
                    b.free()
                    if streamed and i + 1 < k:
                        t2 = time.perf_counter()
                        nxt = make()
                        t_ingest += time.perf_counter() - t2
                return st, w

            run(1)
            t_ingest = t_step = 0.0
            del step_ms[:]
            self.barrier()
            t0 = time.perf_counter()
            st, w = run(e2e_steps)
            self.barrier()
            e2e_s = self.max_over_ranks(time.perf_counter() - t0)
            assert st["batch"] >= n and np.all(np.isfinite(w))
            e2e = {"value": world * n * iters * e2e_steps / e2e_s, "unit": UNIT,
                   "h2d_bytes_per_step": int(world * h2d),
                   **({"host_input_bytes_per_step": int(world * host_input)} if tweets is not None else {}),
                   "d2h_bytes_per_step": int(world * (d * 8 + iters * 8 + 80)),
                   "steps": e2e_steps, "ingest_ms_per_step": t_ingest / e2e_steps * 1e3,
                   "compute_ms_per_step": t_step / e2e_steps * 1e3,
                   "step_ms": [round(x, 3) for x in step_ms[:64]],   # this rank's wall clock per step (ingest call + step)
                   "h2d_GBps_effective_per_gpu": h2d * e2e_steps / e2e_s / 1e9,
                   "overlap": "none: HBM holds one batch of this size, ingest and compute alternate" if streamed else
                              ("none (--e2e-sync)" if self.args.e2e_sync else "step t+1's ingest runs beside step t's compute"),
                   "numa": self.numa,
                   "includes": "host buffers -> HBM ingest (%s), stats + %d iterations, weights/loss/stats back" % (how, iters)}

        narrow = csr and d <= 2048 and not os.environ.get("TWTML_NO_NARROW")  # csr_narrow.cuh: kNarrowMaxD
        rec = None
        if rank == 0:
            peak, peak_src = measured_peak()
            achieved = kbytes / (kern_ms_per_launch * 1e-3) / 1e9
            traffic, traffic_src = traffic_entry(n, d, csr) if world == 1 else (None, None)
            resident_gb = (nnz * 12 * (1 if narrow else 2) + n * 24) / 1e9 if csr else n * (d + 2) * 8 / 1e9
            cfg = workload_config(name, dict(wl, rows=wl["rows"]), world)
            rec = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
                   "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": self.args.scaling, "vs_baseline": None,
                   "dtype": "f64", "data": "synthetic", "config": cfg, "weights_sha256_16": weights_sha,
                   "notes": {"rows_trained_per_gpu": n,
                             **({"tweets_per_gpu": n_tweets, "nnz_per_gpu": nnz} if tweets is not None else {}),
                             "exchange": "one exchange of the shard partials per iteration",
                             "l2": ("inputs (%.3f GB/GPU) about the size of L2, fetched with an evict-first policy: ncu shows every launch "
                                    "reading them from HBM (profiles/); no flush" % resident_gb) if narrow else
                                   "inputs (%.2f GB/GPU) larger than L2; no flush needed" % resident_gb,
                             "timing": "CUDA events on the library stream around each step's iteration loop, max over ranks"},
                   "wall_ms_per_step": wall_ms / steps,
                   "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                                "kernel": ("k_csr_narrow<TRAIN> (single pass over the CSR)" if narrow else "k_csr_rows + k_csc_windows + k_csc_edges (one read of the CSR credited)") if csr else ("k_dense_iter<1,TRAIN,RL1>" if d + 1 > 992 else "k_dense_iter<1,TRAIN,row lanes>"),
                                "kernel_ms": kern_ms_per_launch, "bytes_per_launch": int(kbytes),
                                "frac_of_nominal_8TBs": achieved / 8000.0},
                   "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
                   "exchange": ("none (1 GPU)" if world == 1 else (("NCCL all-to-all of gradient slices + all-gather of w, %d calls per step" if nccl_calls == 2 * iters else
                                                                     "NCCL all-gather, %d calls per step") % nccl_calls if nccl_calls else
                                                                   "peer-memory stores + flags inside the reduce/update kernels (0 NCCL calls per step)"))}
        if batch is not None:
            batch.free()
        model.free()
        return rec

    def close(self):
        self.chunk = None
        self.ctx.close()
        if self.world > 1:
            self.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--sub", default=None, help="comma-separated workloads measured as sub_records (default at N = 1 with the "
                                                 "default workload: %s; 'none' turns them off)" % DEFAULT_SUBS)
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU (testing)")
    ap.add_argument("--iters", type=int, default=0, help="override SGD iterations per micro-batch (testing)")
    ap.add_argument("--e2e-steps", type=int, default=-1, help="timed end-to-end steps (default: 2 for streamed 135 GB batches, else 8)")
    ap.add_argument("--e2e-sync", action="store_true", help="blocking ingest in the e2e leg (A/B against the pipelined default)")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not move the process to the GPU's NUMA node before allocating host buffers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="label of the line: 'strong' when --rows was set to total_rows / N so that the job's work is fixed")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.rows:
        wl["rows"] = args.rows
    if args.iters:
        wl["iters"] = args.iters
    if args.impl == "reference":
        return run_reference(args, args.workload, wl)
    if args.warmup < 3:
        args.warmup = 3
    full_affinity = os.sched_getaffinity(0)

    bench = Bench(args)
    e2e_steps = args.e2e_steps if args.e2e_steps >= 0 else (2 if (not wl.get("csr") and wl["rows"] > 4 * CHUNK_ROWS) else 8)
    line = bench.run(args.workload, wl, args.steps, args.warmup, e2e_steps, sample_clocks=True)
    subs = args.sub if args.sub is not None else (DEFAULT_SUBS if (args.gpus == 1 and args.workload == DEFAULT_WORKLOAD and not args.rows) else "none")
    sub_records = {}
    if subs != "none":
        for name in [s for s in subs.split(",") if s and s != args.workload]:
            swl = dict(WORKLOADS[name])
            rec = bench.run(name, swl, 10, 3, 8, sample_clocks=False)
            if rec is not None:
                sub_records[name] = {k: rec[k] for k in ("value", "unit", "steps", "warmup", "ms_per_step", "config", "notes", "roofline", "e2e", "gpu_launches")}
    bench.close()
    if line is not None:
        if sub_records:
            line["sub_records"] = sub_records
        if args.gpus == 1 and not args.no_cpu_baseline:
            os.sched_setaffinity(0, full_affinity)               # the CPU leg uses every core of the box again
            line["cpu_baseline"], _, _ = cpu_leg(wl, 3, 1, max_seconds=40.0)
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
