"""bench.py -- headline benchmark: DeiT-S/16 224^2 bf16 training step, synthetic batch 256/GPU,
data-parallel over N B200s (BASELINE.json configs[1]).  Prints ONE JSON line on rank 0.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference        # the reference-semantics CPU path (oracle port)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "self-attention-experiments-vision_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "train images/sec DeiT-S/16 224^2 bf16"
MODEL = "deit_s_patch16"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sust=p["bf16_tflops_sustained"],
                    src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.stop = index, [], False
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(float(s[0])) for s in self.samples if s and s[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for n, v in zip(names, s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        mx = max((int(float(s[1])) for s in self.samples if len(s) > 1), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_oracle_throughput(batch=8, reps=2, threads=None):
    """images/sec of the fp32 CPU restatement (oracle port), fwd+bwd, DeiT-S/16 224^2."""
    import torch
    from oracle import vit_oracle as O
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    cfg = O.vit_config(MODEL)
    params = O.synthetic_vit_params(cfg, 224, seed=42)
    images, labels = O.synthetic_batch(batch, 224, seed=0)
    best = None
    for i in range(reps + 1):
        t0 = time.perf_counter()
        O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
        dt = time.perf_counter() - t0
        if i > 0:
            best = dt if best is None else min(best, dt)
    return batch / best, threads, best


def run_reference(args):
    """--impl reference: the reference's own CPU path is JAX/Flax, which is not installable in this
    image; per the tier rules the oracle port (line-by-line fp32 restatement) is timed
    instead, with all host threads, each step = a bounded sample (fwd+bwd of `sample` images)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import vit_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    sample = 64
    cfg = O.vit_config(MODEL)
    params = O.synthetic_vit_params(cfg, 224, seed=42)
    images, labels = O.synthetic_batch(sample, 224, seed=0)
    steps = max(1, min(args.steps, 8))
    warm = max(1, min(args.warmup, 1))
    for _ in range(warm):
        O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.vit_loss_and_grads(params, images, labels, cfg, 0.1)
    dt = (time.perf_counter() - t0) / steps
    v = sample / dt
    desc = f"oracle port (torch CPU fp32), fwd+bwd of {sample} images per step, {threads} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "DeiT-S/16 224^2 training step (fwd+bwd), bounded sample",
                   "sample_images": sample},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": desc},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def _event_time(torch, fn, reps=20, warm=3):
    """Average device time (ms) of fn() with CUDA events on the launching stream."""
    for _ in range(warm):
        fn()
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(st)
    for _ in range(reps):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def kernel_rooflines(torch, ops, B, cfg, pk, launches_per_step_ms):
    """The step's heaviest kernels timed ALONE (CUDA events, every launch streams > L2 of fresh
    bytes) against both roofs on SURVEY 8(d)'s accounting: algorithmic FLOPs (1 MAC = 2, backward
    without recomputation) and algorithmic bytes (every bf16 operand / result once).  The entry
    with the largest share of the step (duration x launches per step / step time) comes first."""
    N, C, H = 197, cfg["embed_dim"], cfg["num_heads"]
    d, hid, M, L = C // H, 4 * C, B * 197, cfg["num_layers"]
    dev = "cuda"
    ents = []

    def entry(name, what, ms, flops, alg_bytes, launches, note=None, moved_bytes=None, ncu_key=None):
        t_tensor = flops / (pk["tf_burst"] * 1e12) * 1e3
        t_hbm = alg_bytes / (pk["hbm"] * 1e9) * 1e3
        bound = "tensor" if t_tensor >= t_hbm else "hbm"
        e = {"kernel": name, "replaces": what, "ms": ms, "launches_per_step": launches,
             "share_of_step": ms * launches / launches_per_step_ms,
             "flops": flops, "algorithmic_bytes": alg_bytes,
             "tflops": flops / (ms * 1e-3) / 1e12, "gbs": alg_bytes / (ms * 1e-3) / 1e9,
             "tensor_floor_ms": t_tensor, "hbm_floor_ms": t_hbm, "bound": bound,
             "frac_of_tensor_peak": flops / (ms * 1e-3) / 1e12 / pk["tf_burst"],
             "frac_of_hbm_peak": alg_bytes / (ms * 1e-3) / 1e9 / pk["hbm"]}
        e["frac"] = e["frac_of_tensor_peak"] if bound == "tensor" else e["frac_of_hbm_peak"]
        if moved_bytes is not None:
            # what the kernel has to move by design (fp32 residual stream, saved pre-activation):
            # the bandwidth it actually sustains, next to the 8(d) figure above
            e["moved_bytes"] = moved_bytes
            e["frac_of_hbm_peak_moved"] = moved_bytes / (ms * 1e-3) / 1e9 / pk["hbm"]
        if note:
            e["note"] = note
        e["ncu_key"] = ncu_key or name.split("<")[0].split(" ")[0]
        ents.append(e)

    # ---- fused attention (attention.py:39-58 and its reverse mode), one encoder layer
    qkv = torch.randn(B, N, 3, H, d, device=dev).bfloat16()
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    o2, lse = ops.attn_fwd(q, k, v)
    d_o = torch.randn(B, N, H, d, device=dev).bfloat16()
    dqkv = torch.empty_like(qkv)
    tok = B * N * H * d * 2           # bytes of one bf16 [B,N,H,d] tensor
    ms = _event_time(torch, lambda: ops.attn_fwd(q, k, v))
    entry("attn_fwd2_kernel", "attention.py:39-58", ms, 4.0 * N * N * d * B * H, 4 * tok, L,
          "q,k,v in, o out; the kernel also writes o_lo (bf16 remainder of O) and lse")
    ms = _event_time(torch, lambda: ops.attn_bwd(q, k, v, o2, d_o, lse, dqkv[:, :, 0], dqkv[:, :, 1],
                                                 dqkv[:, :, 2]))
    entry("attn_bwd3_kernel", "reverse mode of attention.py:39-58", ms, 8.0 * N * N * d * B * H, 8 * tok, L,
          "q,k,v,o,dO in, dq,dk,dv out (10 N^2 d executed with S recomputed; also reads o_lo)")
    del qkv, o2, d_o, dqkv
    # ---- FC1 GEMM, bias + GELU epilogue (ff.py:24-25)
    x = torch.randn(M, C, device=dev).bfloat16()
    w = (torch.randn(C, hid, device=dev) / C ** 0.5).bfloat16()
    b = torch.randn(hid, device=dev)
    out = torch.empty(M, hid, device=dev, dtype=torch.bfloat16)
    pre = torch.empty_like(out)
    ms = _event_time(torch, lambda: ops.gemm(x, w, b_mn=True, out=out, bias=b, act=1, aux_out=pre))
    entry("gemm_kernel<256,K,MN,pair> FC1", "ff.py:24-25", ms, 2.0 * M * C * hid,
          M * C * 2 + C * hid * 2 + hid * 4 + M * hid * 2, L,
          "8(d) accounting: ONE bf16 output; the kernel writes two (pre-activation saved for the "
          "backward + activation) against a measured write-only HBM ceiling of 3.9 TB/s",
          moved_bytes=M * C * 2 + C * hid * 2 + hid * 4 + 2 * M * hid * 2)
    del out, pre, w, b
    # ---- FC2 GEMM with bias + residual + the following LayerNorm in its epilogue (ff.py:27, vit.py:31,19)
    h = torch.randn(M, hid, device=dev).bfloat16()
    w2 = (torch.randn(hid, C, device=dev) / hid ** 0.5).bfloat16()
    b2 = torch.randn(C, device=dev)
    res = torch.randn(M, C, device=dev)
    out32 = torch.empty(M, C, device=dev)
    g1, be1 = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    if ops.gemm_ln_tiles(C):
        ms = _event_time(torch, lambda: ops.gemm(h, w2, b_mn=True, out=out32, bias=b2, residual=res,
                                                 ln=(g1, be1, 1e-6)))
        entry("gemm_kernel<128,K,MN,pair> FC2+LN", "ff.py:27 + vit.py:31 + vit.py:19", ms, 2.0 * M * C * hid,
              M * hid * 2 + C * hid * 2 + 2 * M * C * 2 + M * C * 2, L,
              "8(d) accounting: bf16 stream in / out + bf16 LayerNorm output; the stream is fp32 here",
              moved_bytes=M * hid * 2 + C * hid * 2 + 2 * M * C * 4 + M * C * 2, ncu_key="gemm_kernel_fc2_fwd_ln")
    del h, w2, res, out32
    # ---- LayerNorm backward (+ stream-gradient fan-in, bf16 copy, bias column sums)
    xs = torch.randn(M, C, device=dev)
    dy = torch.randn(M, C, device=dev).bfloat16()
    g_in = torch.randn(M, C, device=dev)
    dx = torch.empty(M, C, device=dev)
    dx16 = torch.empty(M, C, device=dev, dtype=torch.bfloat16)
    sc, bi = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    dsc, dbi = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    _y, mean, rstd = ops.layernorm_fwd(xs, sc, bi)
    ms = _event_time(torch, lambda: ops.layernorm_bwd(dy, xs, sc, mean, rstd, dsc, dbi, dx_add=g_in, dx=dx,
                                                      dx16=dx16))
    entry("ln_bwd_pieces_kernel", "reverse mode of nn.LayerNorm (vit.py:19,26,57)", ms, 0.0, 3 * M * C * 2,
          2 * L + 1, "8(d): 3 x bf16 per element; the kernel moves 16 B per element (fp32 x, fp32 stream "
                     "gradient in and out, bf16 dy, bf16 copy)", moved_bytes=16 * M * C)
    ents.sort(key=lambda e: -e["share_of_step"])
    return ents


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=256, help="images per GPU")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import sav
    from models import create_model
    from oracle import vit_oracle as O  # synthetic inputs + cpu_baseline leg only
    from sav import _abi, ops
    from sav.train import GraphedTrainStep, HWCN, TrainState, train_step

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    _abi.check(_abi.lib().sav_device_check())
    dp = None
    if world > 1:
        from sav.dp import DataParallel
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # development knobs for the exchange: bucket size (MB) and overlap with backward
        dp = DataParallel(bucket_mb=float(os.environ.get("SAV_DP_BUCKET_MB", "16")),
                          overlap=os.environ.get("SAV_DP_OVERLAP", "1") == "1")
    W = max(3, args.warmup)
    K = args.steps
    B = args.batch
    cfg = O.vit_config(MODEL)

    model = create_model(MODEL)
    images_host, labels_host = O.synthetic_batch(B, 224, seed=rank)
    images_host, labels_host = images_host.pin_memory(), labels_host.pin_memory()
    images = images_host.cuda(non_blocking=True)
    labels = labels_host.cuda(non_blocking=True)
    variables = model.init({"params": 42}, HWCN(images), is_training=True)
    # the reference's zero-init head (vit.py:98) makes every other gradient exactly 0; use the
    # synthetic non-degenerate parameters of SURVEY 8d so the step does real work
    variables = {"params": variables["params"].arena.load_tree(O.synthetic_vit_params(cfg, 224, seed=42))}
    if dp is not None:
        dp.broadcast_params(variables["params"].arena)
    state = TrainState(model, variables, lr=5e-4 * B * world / 512, weight_decay=1e-4, clip_grad=1.0)

    x = HWCN(images)
    if args.no_graph:
        def step():
            state.advance_schedule()
            return train_step(state, x, labels, 0.1, dp)
        n0 = _abi.launch_count(); step(); launches = _abi.launch_count() - n0
    else:
        g = GraphedTrainStep(state, x, labels, 0.1, dp)
        step, launches = g, g.launches_per_step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        st = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(st)
        for _ in range(n):
            fn()
        e1.record(st)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(W):
        loss = step()
    clk = ClockSampler(local)
    clk.__enter__()
    total_ms = timed(step, K)
    loss_val = loss.item()
    value = B * world * K / (total_ms * 1e-3)

    # e2e: every step copies its inputs from pinned host memory and reads the loss back.  Through
    # the graphed step's host-fed API the copy of batch i+1 is issued right after step i is
    # launched (double-buffered staging, sav/train.py), so K host->device copies happen inside
    # the timed region and only the first one is exposed.
    if args.no_graph:
        def e2e_step():
            images.copy_(images_host, non_blocking=True)
            labels.copy_(labels_host, non_blocking=True)
            l = step()
            return l.item()

        def e2e_loop(n):
            for _ in range(n):
                e2e_step()
    else:
        def e2e_loop(n):
            step.prefetch(images_host, labels_host)
            for i in range(n):
                l = step.step_prefetched()
                if i + 1 < n:
                    step.prefetch(images_host, labels_host)
                l.item()
    e2e_loop(2)
    e2e_ms = timed(lambda: e2e_loop(K), 1)
    clk.__exit__()   # clocks / throttle reasons sampled across both timed regions
    e2e_value = B * world * K / (e2e_ms * 1e-3)

    pk = peaks()
    out = {
        "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "DeiT-S/16 224^2 training step (fwd+bwd+grad mean+clip+Adam)",
                   "batch_per_gpu": B, "global_batch": B * world, "tokens": 197,
                   "parallelism": f"dp{world}", "cuda_graph": not args.no_graph,
                   "grad_exchange": (dp.stats.get("transport", dp.transport) if dp is not None else None),
                   "l2": "inputs larger than L2 (activations ~GBs per step)"},
        "e2e": {"value": e2e_value, "unit": "images/s",
                "h2d_bytes_per_step": images_host.numel() * 4 + labels_host.numel() * 4,
                "d2h_bytes_per_step": 8},
        "gpu_launches": launches * K,
        "clocks": clk.summary(),
        "loss": loss_val,
    }
    if rank == 0:
        fl = 3 * 9.198e9 * B  # train FLOP/img = 3 x fwd
        out["step_tflops"] = fl / (total_ms / K * 1e-3) / 1e12
        out["step_frac_of_sustained_peak"] = out["step_tflops"] / pk["tf_sust"]
        # Roofline block: the heaviest kernels of the step timed alone; `roofline` is the one with
        # the largest share of the step, `kernels` lists all of them with both roofs.  BASELINE's
        # metric names "attn TC-pipe % of peak": attention_tc_pipe carries FLOP/time against the
        # measured dense-bf16 burst peak for both attention kernels and the ncu tensor-pipe counter
        # of the same shapes (profiles/, captured under ncu, not in this run).
        ents = kernel_rooflines(torch, ops, B, cfg, pk, total_ms / K)
        top = ents[0]
        try:
            with open(os.path.join(ROOT, "profiles", "r02_ncu_counters.json")) as f:
                ncu = json.load(f)
        except Exception:
            ncu = {}
        for e in ents:
            c = ncu.get(e.pop("ncu_key"), {})
            e["traffic"] = c.get("dram_bytes")
            if "tensor_pipe_pct" in c:
                e["ncu_tensor_pipe_pct"] = c["tensor_pipe_pct"]
        out["roofline"] = {"bound": top["bound"], "kernel": top["kernel"],
                           "achieved": top["gbs"] if top["bound"] == "hbm" else top["tflops"],
                           "peak": pk["hbm"] if top["bound"] == "hbm" else pk["tf_burst"],
                           "unit": "GB/s" if top["bound"] == "hbm" else "TFLOP/s",
                           "frac": top["frac"], "traffic": top.get("traffic"),
                           "moved_bytes": top.get("moved_bytes"),
                           "frac_on_moved_bytes": top.get("frac_of_hbm_peak_moved"),
                           "algorithmic_bytes": top["algorithmic_bytes"], "flops": top["flops"],
                           "ms": top["ms"], "share_of_step": top["share_of_step"],
                           "peak_source": pk["src"] + " (burst / copy figures; kernel timed alone)"}
        out["kernels"] = ents
        att = {e["kernel"]: e for e in ents if e["kernel"].startswith("attn")}
        out["attention_tc_pipe"] = {
            k: {"tflops": e["tflops"], "pct_of_dense_bf16_peak": 100.0 * e["frac_of_tensor_peak"],
                "ncu_tensor_pipe_pct": e.get("ncu_tensor_pipe_pct")} for k, e in att.items()}
        if world == 1 and not args.no_cpu_baseline:
            v, threads, dt = cpu_oracle_throughput(batch=64, reps=5)
            out["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": threads, "kind": "port",
                                   "sample": f"fwd+bwd of 64 images (of the 256-image step), best of 5 "
                                             f"after 1 warm-up, DeiT-S/16 224^2, fp32 torch-CPU "
                                             f"oracle, {dt:.1f} s per pass; reference JAX CPU path "
                                             f"not installable in this image"}
        print(json.dumps(out), flush=True)
    if world > 1:
        # Clean teardown: the captured step holds NCCL kernels, so the graph (and everything that
        # refers to it) goes first, then the communicator.  destroy_process_group() runs under a
        # watchdog -- if a rank is still inside NCCL after 20 s the bench has printed its line long
        # ago and the process leaves through os._exit rather than hang the driver.
        import gc
        import threading
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        step = g = state = dp = None   # noqa: F841
        gc.collect()
        torch.cuda.synchronize()
        done = threading.Event()

        def _teardown():
            try:
                dist.destroy_process_group()
            finally:
                done.set()
        t = threading.Thread(target=_teardown, daemon=True)
        t.start()
        if not done.wait(20.0):
            sys.stderr.write("bench: communicator teardown did not finish in 20 s, leaving\n")
            sys.stderr.flush()
            os._exit(0)


if __name__ == "__main__":
    main()
