#!/usr/bin/env python
"""Headline benchmark: training samples/s (forward + backward) of the VISITRON encoder hot path.

    python bench.py --gpus N --steps K --warmup W            # our arm (N > 1: launched by torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W   # reference arm: the CPU port on the host cores

One step = one pass of the hot path over one synthetic batch: PreTrainOscar forward (encoder over
[511 dialog tokens ; 256 region slots], masked-LM / detector-class / next-action heads, loss) and
backward (all 113 M parameter gradients, averaged across ranks by the bucketed all-reduce that
overlaps backward).  The optimizer is excluded, as in BASELINE.json's metric (fwd+bwd).

Prints ONE JSON line (rank 0): `value` = whole-job samples/s with the batch resident in HBM,
`e2e` = the same through the public API from pinned HOST buffers (H2D of the step's batch and a D2H
read of the loss inside the timed region), `roofline` for the dominant kernel (the tcgen05 GEMM),
`cpu_baseline` = the oracle port timed on this box's host cores (N = 1 only).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "train samples/sec (fwd+bwd)"
UNIT = "samples/s"
H, I, V, DC, A, IMG, LOC = 768, 3072, 30522, 1601, 36, 2054, 128


def flops_per_sample(S, R, layers=12, all_rows_heads=True, labelled=(0, 0)):
    """Forward FLOPs of one sample; training = 3x."""
    enc_gemm = layers * 24 * S * H * H
    attn = layers * 4 * S * S * H
    img = 2 * R * (IMG + LOC) * H
    pool = 2 * H * H
    if all_rows_heads:
        heads = S * (2 * H * H + 2 * H * V) + 2 * S * H * DC + 2 * H * A
    else:
        heads = labelled[0] * (2 * H * H + 2 * H * V) + 2 * labelled[1] * H * DC + 2 * H * A
    return dict(encoder_gemm=enc_gemm, attention=attn, img=img, pooler=pool, heads=heads,
                total=enc_gemm + attn + img + pool + heads)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(tflops=d.get("bf16_tflops_sustained", d.get("bf16_tflops")), burst=d.get("bf16_tflops"), hbm=d.get("hbm_gbs"),
                    src="measured (MEASURED_PEAKS.json, sustained)")
    return dict(tflops=1400.0, burst=1590.0, hbm=6650.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    """SM clock / power / throttle reasons DURING the timed region, sampled in-process through NVML every 50 ms.
    (A `nvidia-smi -lms` subprocess started at the top of the timed region was tried first: its start-up and queries contend
    for the driver with the kernel launches and slowed the region it was observing by 5-300 %.)"""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        self.rows, self.on, self.alive, self.h, self.nv = [], False, False, None, None
        try:
            import pynvml as nv
            nv.nvmlInit()
            pr = torch.cuda.get_device_properties(index)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            self.h, self.nv = nv.nvmlDeviceGetHandleByPciBusId(bus.encode()), nv
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.alive = True
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        except Exception as e:   # no NVML: say so in the JSON line
            self.err = "%s: %s" % (type(e).__name__, e)

    def _one(self):
        nv = self.nv
        try:
            mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        return (float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)), nv.nvmlDeviceGetPowerUsage(self.h) / 1e3, int(mask), time.perf_counter())

    def _run(self):
        while self.alive:
            if self.on:
                try:
                    self.rows.append(self._one())
                except Exception:
                    pass
            time.sleep(0.05)

    def start(self):
        self.rows, self.on = [], True

    def window(self, t0, t1):
        """median SM clock / max power of the samples taken between two host times"""
        rows = [r for r in list(self.rows) if t0 <= r[3] <= t1]
        return (statistics.median(r[0] for r in rows), round(max(r[1] for r in rows), 1)) if rows else (None, None)

    def stop(self):
        self.on = False
        if self.h is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["NVML unavailable (%s)" % getattr(self, "err", "?")])
        rows = list(self.rows)
        reasons = sorted({name for _, _, m, _ in rows for bit, name in self.REASONS if m & bit})
        sm = [r[0] for r in rows]
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=self.max_mhz, power_w_max=max((r[1] for r in rows), default=None),
                    samples=len(rows), reasons=reasons)


def build_model(args, device):
    from visitron_b200 import modeling
    from visitron_b200.pretrained import BertConfig
    cfg = BertConfig(V, hidden_dropout_prob=args.dropout, attention_probs_dropout_prob=0.1 if args.dropout > 0 else 0.0)
    cfg.img_feature_dim, cfg.detector_classes, cfg.action_space = IMG, DC, A
    torch.manual_seed(88)
    if args.workload != "pretrain":
        model = modeling.BertImgModelwithLocationEmbeds(cfg)   # what train.py hands the agent as `model.bert` (vs/train.py:47)
    else:
        model = modeling.PreTrainOscar(cfg)     # random-init bert-base (hf/modeling_bert.py:550-561), no checkpoint
        model.vt_packed = not args.dense
    return model.to(device).train()


VS_STEPS, VS_CAND, VS_FEAT, VS_HID = 10, 12, 2052, 512   # decoder steps, candidates incl. STOP, 2048+4 features, LSTM width
AGENT_DROPOUT = 0.5                                      # --dropout of the agents' modules (vs/params.py:300, every run script)
TB_STEPS = 40                                            # turn_based episode length (run_scripts/turn_based_train/ndh_rxr.sh)


class Workload:
    """One step of the hot path for `--workload`: host batches, the step function and the algorithmic FLOPs."""

    def __init__(self, args, device, rank, world):
        from visitron_b200 import agent_heads, agent_models, synthetic
        self.args, self.device, self.world = args, device, world
        self.model = build_model(args, device)
        b, n_host = args.batch, 4
        self.extra_params = []
        self.flop_mult = 3.0   # forward + backward
        self.S, self.R = 511, 0
        self.labelled = (0.0, 0.0)
        self.trace = None
        wk = args.workload
        if wk in ("viewpoint_select", "turn_based"):
            # BASELINE configs[1] / configs[3]: one fine-tuning iteration of the reference's agent (vs/agent.py:493-518, tb/agent.py:
            # 405-430) -- OscarEncoder (encoder at S = 511 text-only with token types and the ~uint8 mask, the 511-step LSTM, the
            # decoder's initial state), then the rollout of AttnDecoderLSTM over a recorded trace on the device (10 steps of
            # feature attention over 36 views + LSTMCell + context attention over 511 + candidate scorer over 12 candidates; resp.
            # 40 steps of action embedding + LSTMCell + context attention + 6-way logits), teacher forcing, masked CE per step,
            # loss / episode_len, backward through everything.  Dropout: the encoder's --drop_out, the agent modules' 0.5.
            torch.manual_seed(88)
            steps = VS_STEPS if wk == "viewpoint_select" else TB_STEPS
            self.steps = steps
            self.encoder = agent_models.OscarEncoder(None, self.model, VS_HID, VS_HID, AGENT_DROPOUT if args.dropout > 0 else 0.0).to(device).train()
            if wk == "viewpoint_select":
                self.decoder = agent_models.AttnDecoderLSTM(4, 64, VS_HID, AGENT_DROPOUT if args.dropout > 0 else 0.0, VS_FEAT).to(device).train()
                mk = lambda i: synthetic.viewpoint_trace(steps, b, seed=88 + rank + 1000 * i, candidates=VS_CAND, feat=VS_FEAT)  # noqa: E731
                dec_flops = (2 * 4 * 64 + 2 * VS_HID * VS_FEAT + 4 * 36 * VS_FEAT + 2 * (64 + VS_FEAT + VS_HID) * 4 * VS_HID
                             + 2 * VS_HID * VS_HID + 4 * 511 * VS_HID + 4 * VS_HID * VS_HID + 2 * VS_HID * VS_FEAT + 2 * VS_CAND * VS_FEAT)
            else:
                self.decoder = agent_models.TurnAttnDecoderLSTM(8, 6, 32, VS_HID, AGENT_DROPOUT if args.dropout > 0 else 0.0, 2048).to(device).train()
                mk = lambda i: synthetic.turn_trace(steps, b, seed=88 + rank + 1000 * i)  # noqa: E731
                dec_flops = (2 * (32 + 2048 + VS_HID) * 4 * VS_HID + 2 * VS_HID * VS_HID + 4 * 511 * VS_HID + 4 * VS_HID * VS_HID + 2 * 6 * VS_HID)
            for m in (self.encoder, self.decoder):      # nn.LSTM / nn.Linear default init
                pass
            self.extra_params = [p for n, p in self.encoder.named_parameters() if not n.startswith("bert.")] + list(self.decoder.parameters())
            host = []
            for i in range(n_host):
                hb = synthetic.agent_batch(b, seed=88 + rank + 1000 * i)     # the scripts' situation: every length 511, mask all zero (F6)
                hb.update(mk(i))
                host.append(hb)
            lstm_flops = 511 * 2 * (768 + VS_HID) * 4 * VS_HID
            self.heads_flops = lstm_flops + steps * dec_flops
        elif wk == "qa_classifier":
            # BASELINE configs[4]: the encoder forward-only under no_grad + question_linear / masked BCE forward + backward
            # (vs/classifier/agent.py:158-163,493-503), swept over --batch.
            g = torch.Generator().manual_seed(88)
            mk = lambda *s: torch.nn.Parameter((0.05 * torch.randn(*s, generator=g)).to(device))  # noqa: E731
            self.qa = dict(w1=mk(VS_HID // 2, VS_HID), b1=mk(VS_HID // 2), w2=mk(1, VS_HID // 2), b2=mk(1))
            self.extra_params = list(self.qa.values())
            host = []
            for i in range(n_host):
                hb = synthetic.finetune_encoder_batch(b, seed=88 + rank + 1000 * i, valid_len=511)
                gi = torch.Generator().manual_seed(88 + rank + 1000 * i)
                hb["target"] = (torch.rand(b, generator=gi) < 0.4).float()
                hb["valid"] = (torch.rand(b, generator=gi) >= 0.25).to(torch.uint8)
                host.append(hb)
            self.heads_flops = 2 * (VS_HID * VS_HID // 2 + VS_HID // 2)
            self.flop_mult = 1.0
        else:
            host = [synthetic.pretrain_batch(b, seed=88 + rank + 1000 * i) for i in range(n_host)]  # per-rank shard, seed 88+rank
            if args.feat_bf16:   # a bf16 feature cache: half the H2D bytes; the default is the loader's fp32
                for hb in host:
                    hb["img_feats"], hb["img_location_embeddings"] = hb["img_feats"].bfloat16(), hb["img_location_embeddings"].bfloat16()
            self.S, self.R = 767, 256
            self.labelled = (sum(int((hb["labels"] != -1).sum()) for hb in host) / (n_host * b),
                             sum(int((hb["token_labels"] != -1).sum()) for hb in host) / (n_host * b))
            self.heads_flops = None
            # padding-free execution: the encoder runs on the rows that exist (mask != 0, labelled, or position 0)
            keep = [((hb["attention_mask"] != 0) | (hb["labels"] != -1) | (hb["token_labels"] != -1)) for hb in host]
            for k in keep:
                k[:, 0] = True
            lens = torch.cat([k.sum(1) for k in keep]).double()
            self.valid_rows = (float(lens.mean()), float((lens * lens).mean()))     # E[S_valid], E[S_valid^2] per sample
        self.host = [{k: v.pin_memory() for k, v in hb.items()} for hb in host]
        self.n_host = n_host
        self.params = list(self.model.parameters()) + self.extra_params
        self._heads = agent_heads

    def forward(self, batch):
        """The step's forward: a tuple of 0-d tensors, the loss first (what GraphedStep captures and the eager step runs)."""
        wk = self.args.workload
        if wk == "pretrain":
            return self.model(**batch)
        if wk == "qa_classifier":
            with torch.no_grad():   # vs/classifier: the encoder is frozen
                seq, _ = self.model(batch["input_ids"], token_type_ids=batch["token_type_ids"], attention_mask=batch["attention_mask"])
            q = self.qa
            _, loss = self._heads.qa_head(q["w1"], q["b1"], q["w2"], q["b2"], seq[:, 0, :VS_HID], batch["target"], batch["valid"])
            return (loss,)
        from visitron_b200 import rollout
        enc_batch = dict(input_ids=batch["input_ids"], token_type_ids=batch["token_type_ids"], mask=batch["mask"], lengths=batch["lengths"])
        if wk == "turn_based":
            loss, _ = rollout.turn_based_rollout(self.encoder, self.decoder, enc_batch, batch, self.steps)
        else:
            loss, _ = rollout.viewpoint_rollout(self.encoder, self.decoder, enc_batch, batch, self.steps)
        return (loss,)

    def post_backward(self):
        """The head parameters live outside the encoder's flat buffer: one small all-reduce each (pre-training has none)."""
        if self.world > 1:
            for p in self.extra_params:
                if p.grad is not None:
                    torch.distributed.all_reduce(p.grad, op=torch.distributed.ReduceOp.AVG)

    def step(self, batch):
        for p in self.params:
            p.grad = None
        out = self.forward(batch)
        out[0].backward()
        self.post_backward()
        return out

    def flops(self, executed=False, valid_only=False):
        fl = flops_per_sample(self.S, self.R, all_rows_heads=not executed, labelled=self.labelled)
        if valid_only and getattr(self, "valid_rows", None):   # the encoder on the rows that exist: GEMMs ~ E[S_valid], attention ~ E[S_valid^2]
            sv, sv2 = self.valid_rows
            fl["total"] -= fl["encoder_gemm"] + fl["attention"]
            fl["encoder_gemm"] = 12 * 24 * sv * H * H
            fl["attention"] = 12 * 4 * sv2 * H
            fl["total"] += fl["encoder_gemm"] + fl["attention"]
        if self.heads_flops is not None:   # no pre-training heads, no image projection on this workload
            fl["total"] += self.heads_flops - fl["heads"] - fl["img"]
            fl["heads"], fl["img"] = self.heads_flops, 0
        return fl


def workload_desc(args):
    wk = args.workload
    if wk == "viewpoint_select":
        return ("viewpoint_select_step: one fine-tuning iteration of the viewpoint-select agent (vs/agent.py:493-518): OscarEncoder (encoder "
                "fwd+bwd at T=511 text-only with token types and the ~uint8 mask, nn.LSTM 768->512 over the 511 positions, decoder init) + "
                "AttnDecoderLSTM rollout over a device-resident trace, 10 steps (attention over 36 views x 2052, LSTMCell, attention over the "
                "511 context positions, SoftDotAttention candidate scorer over 12 candidates), teacher forcing, masked CE, bert-base random "
                "init [BASELINE configs[1]]"), 511
    if wk == "turn_based":
        return ("turn_based_step: one fine-tuning iteration of the turn-based agent (tb/agent.py:405-430) at the full RxR text length "
                "(T=511): OscarEncoder (encoder fwd+bwd, nn.LSTM 768->512 over 511 positions) + AttnDecoderLSTM rollout over a "
                "device-resident trace, 40 steps (action embedding, LSTMCell over [32 ; 2048], attention over the 511 context positions, "
                "6-way logits, forward masked, CE ignore 7), teacher forcing, bert-base random init [BASELINE configs[3]]"), 511
    if wk == "qa_classifier":
        return ("qa_classifier_step: encoder forward-only under no_grad (T=511) + question_linear (512->256->1, tanh) with "
                "BCEWithLogits(pos_weight 5) masked mean, forward + backward [BASELINE configs[4]: sweep --batch]"), 511
    return ("pretrain_step: PreTrainOscar fwd+bwd, bert-base random init, NDH-shaped batch (T=511 text + R=256 region slots, "
            "img_feature_dim 2054, 15% MLM + masked detector-class + 36-way action labels), all objectives [BASELINE configs[2]; "
            "configs[1] = --workload viewpoint_select]"), 767




def make_config(args, world):
    """The `config` object of the JSON line: shared by both arms (the reference arm times bounded samples of THIS workload)."""
    desc, S = workload_desc(args)
    b = args.batch
    layout = {}
    if args.workload == "pretrain":
        layout = dict(layout="dense [B, S] rows (padding included)" if args.dense else
                      "packed rows: padding-free execution, the rows that are masked out and unlabelled are dropped on the device")
    return dict(workload=desc, per_gpu_batch=b, global_batch=b * world, seq_len=S, parallelism="dp%d" % world, dropout=args.dropout, **layout,
                l2="per-step working set (%.1f GB of activations) >> 126 MB L2; 4 rotating batches" % (b * S * 24.5e3 * 12 / 1e9),
                grad_allreduce="fp32, one bucket per encoder layer, overlapped with backward" if world > 1 else "none (1 GPU)")


def run_ours(args):
    from visitron_b200 import lib, parallel
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun (one process per GPU)" % args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    saved_stdout = os.dup(1)
    sys.stdout.flush()
    os.dup2(2, 1)   # NCCL prints its version banner on stdout: stdout carries ONE JSON line, everything else goes to stderr
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=device)
    wl = Workload(args, device, rank, world)
    model = wl.model
    if world > 1:
        parallel.distribute(model)
    b = args.batch
    n_host, host, step = wl.n_host, wl.host, wl.step
    dev_batches = [{k: v.to(device) for k, v in hb.items()} for hb in host]
    labelled = wl.labelled

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t)
        return ms

    # warm-up: at least 3 steps AND every rotating batch once (their labelled-row counts differ, so a batch seen for the first
    # time inside the timed region would pay the caching allocator's cudaMalloc there)
    # ... and for at least --min-warm-s of wall time: the first ~0.2 s after a cold start run below the steady clocks (a timed
    # region right behind 3 steps measured 10-25 % low against the same region a second later)
    sampler = ClockSampler(local) if (rank == 0 and not args.no_clocks) else None   # NVML is initialised here, well before the timed region
    n_warm, t0 = 0, time.perf_counter()
    while n_warm < max(args.warmup, 3, n_host + 1) or time.perf_counter() - t0 < args.min_warm_s:
        step(dev_batches[n_warm % n_host])
        torch.cuda.synchronize()
        n_warm += 1
        if n_warm == 1:
            t0 = time.perf_counter()   # the first step carries module loading and allocator growth: the clock starts after it
    # ---- the step as ONE CUDA graph (visitron_b200/graph.py): fixed launch sequence, no host sync, one cudaGraphLaunch per step ----
    graphed, graph_err = None, None
    if not args.no_graph:
        from visitron_b200 import graph as vgraph
        try:
            graphed = vgraph.GraphedStep(model, dev_batches[0], step_fn=lambda m, bt: wl.forward(bt), extra_params=wl.extra_params,
                                         post_backward=wl.post_backward if (world > 1 and wl.extra_params) else None)
        except Exception as e:   # reported in the line, the eager path is timed instead
            graph_err = "%s: %s" % (type(e).__name__, str(e)[:300])
            torch.cuda.synchronize()
        if world > 1:   # every rank must take the same path
            ok = torch.tensor([1.0 if graphed is not None else 0.0], device=device)
            torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)
            if float(ok) == 0.0:
                graphed = None
                graph_err = graph_err or "capture failed on another rank"

    def run_step(bt):
        if graphed is None:
            return torch.stack([o.detach().float() for o in step(bt)])
        graphed.load(bt)
        return graphed.replay()

    for i in range(n_host):   # once more over every batch on the path that is timed
        run_step(dev_batches[i])
    torch.cuda.synchronize()
    # eager comparison point (the public API without the graph): one region
    eager_ms = timed(lambda i: step(dev_batches[i % n_host]), args.steps) / args.steps if graphed is not None else None
    # ---- device-resident throughput (`value`) ----
    if sampler:
        sampler.start()
    # --repeats regions of EXACTLY --steps steps each, every one bracketed as the contract says; `value` is the MEDIAN region and
    # all of them are listed in the line with the SM clock and board power sampled inside each.  Why: about one 0.24 s region in
    # ten is hit by a 100-300 ms stall of the HOST thread on the shared box (profiles/r1_region_diag.txt: the slow regions show
    # HIGHER clocks and LOWER power, i.e. an idle GPU, with and without Python's cyclic GC) -- e.g. 34.9 or 57.9 ms per step
    # between regions of 23.7.  The median of 5 ignores up to two such regions; `e2e` is measured the same way.
    regions, region_clk = [], []
    if args.manual_gc:   # diagnosis: keep the cyclic collector out of the timed regions
        import gc
        gc.collect()
        gc.freeze()
        gc.disable()
    for _ in range(max(1, args.repeats)):
        lib.reset_launch_count()
        h0 = time.perf_counter()
        regions.append(timed(lambda i: run_step(dev_batches[i % n_host]), args.steps))
        launches = lib.launch_count() if graphed is None else graphed.launches_per_step * args.steps
        region_clk.append(sampler.window(h0, time.perf_counter()) if sampler else (None, None))
    ms = statistics.median(regions)
    clocks = sampler.stop() if sampler else None
    value = world * b * args.steps / (ms / 1e3)

    # ---- end to end from pinned host memory through the public API (`e2e`) ----
    from visitron_b200 import input_pipeline
    h2d_bytes = sum(v.numel() * v.element_size() for v in host[0].values())

    def host_stream():   # the loader: pinned host batches, for ever
        i = 0
        while True:
            yield host[i % n_host]
            i += 1

    feed = input_pipeline.DevicePrefetcher(host_stream(), device)   # next batch's H2D overlaps this step's compute

    def e2e_step(i):
        return run_step(next(feed)).cpu()   # D2H read of the step's scalars (host sync, as the loop's .item())

    e2e_step(0)
    e2e_regions = []
    for _ in range(max(1, args.repeats)):
        e2e_regions.append(timed(e2e_step, args.steps))
    e2e_ms = statistics.median(e2e_regions)
    e2e_value = world * b * args.steps / (e2e_ms / 1e3)

    # ---- roofline of the dominant kernel: per-launch CUDA events around every GEMM, same K steps ----
    roof = None
    if not args.no_roofline:
        lib.GEMM_EVENTS = []
        timed(lambda i: step(dev_batches[i % n_host]), args.steps)
        evs, lib.GEMM_EVENTS = lib.GEMM_EVENTS, None
        torch.cuda.synchronize()
        # FLOPs EXECUTED per launch: a GEMM under a device-resident row limit (packed rows, labelled rows) walks min(rows, limit) rows
        # (the limit is the count rounded up to 128 (+128 for the packed layout): those padding rows are computed and are counted here)
        def executed_flops(ev):
            _, _, M_, N_, K_, a_major, lim = ev
            if lim is not None:
                rows_dev = int(lim[0][4 * lim[1]])
                if a_major:
                    K_ = min(K_, rows_dev)
                else:
                    M_ = min(M_, rows_dev)
            return 2.0 * M_ * N_ * K_
        tot_ms = sum(ev[0].elapsed_time(ev[1]) for ev in evs)
        tot_fl = sum(executed_flops(ev) for ev in evs)
        pk = peaks()
        ach = tot_fl / (tot_ms / 1e3) / 1e12 if tot_ms > 0 else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "gemm_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("dram_bytes_per_launch")
        roof = dict(bound="tensor", kernel="gemm_bf16_kernel (tcgen05/TMEM, TMA-fed)", achieved=round(ach, 1), peak=pk["tflops"], unit="TFLOP/s",
                    frac=round(ach / pk["tflops"], 4), traffic=traffic, peak_source=pk["src"], launches=len(evs),
                    gemm_ms_per_step=round(tot_ms / args.steps, 3), gemm_share_of_step=round(tot_ms / args.steps / (ms / args.steps), 3),
                    flops_per_launch_avg=tot_fl / max(len(evs), 1),
                    note="2*M*N*K EXECUTED per launch (M or K = the device-resident row count where a launch runs under a row limit) / CUDA-event "
                         "time per launch, separate pass of the same K steps through the eager path")

    # ---- the step after backward, reported separately (the metric is fwd+bwd): fused AdamW, one launch ----
    from visitron_b200 import optimization as vopt
    no_decay = ["bias", "LayerNorm.weight"]
    named = [(n, p) for n, p in model.named_parameters() if p.grad is not None] + [("head.weight", p) for p in wl.extra_params if p.grad is not None]
    opt = vopt.AdamW([dict(params=[p for n, p in named if not any(nd in n for nd in no_decay)], weight_decay=0.01),
                      dict(params=[p for n, p in named if any(nd in n for nd in no_decay)], weight_decay=0.0)], lr=5e-5, eps=1e-8)
    opt.step()
    opt_ms = timed(lambda i: opt.step(), 5) / 5
    opt_elems = sum(p.numel() for _, p in named)
    # (qa_classifier updates only the 131 k parameters of question_linear: the figure is launch latency there)

    # ---- N > 1: are the exchanged gradients the same on every rank, and are they the mean of the per-rank gradients? ----
    sync_check = grad_sync_check(wl, model, dev_batches[0], device, world) if world > 1 else None

    if rank != 0:
        _leave(world, graphed)
        return
    S = wl.S
    fl, fl_done = wl.flops(), wl.flops(executed=True)
    packed = args.workload == "pretrain" and getattr(model, "vt_packed", False)
    fl_valid = wl.flops(executed=True, valid_only=True) if packed else fl_done
    n_out = 7 if args.workload == "pretrain" else 1
    pk = peaks()
    line = dict(
        metric=METRIC, value=round(value, 2), unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup, warmup_run=n_warm,
        ms_per_step=round(ms / args.steps, 3), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16",
        data="synthetic", timed_regions=dict(n=len(regions), steps_each=args.steps, ms_per_step=[round(r / args.steps, 3) for r in regions],
                                             sm_mhz=[c[0] for c in region_clk], power_w_max=[c[1] for c in region_clk], reported="median"),
        config=make_config(args, world),
        e2e=dict(value=round(e2e_value, 2), unit=UNIT, h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=4 * n_out,
                 ms_per_step=round(e2e_ms / args.steps, 3), regions_ms_per_step=[round(r / args.steps, 3) for r in e2e_regions], reported="median"),
        gpu_launches=int(launches),
        graph=dict(enabled=graphed is not None, launches_per_step=(graphed.launches_per_step if graphed is not None else None), error=graph_err,
                   eager_ms_per_step=(round(eager_ms, 3) if eager_ms is not None else None),
                   note="value / e2e replay ONE captured CUDA graph per step (visitron_b200.graph.GraphedStep: device-side row compaction and "
                        "dropout seed, no host sync); eager_ms_per_step = the same step through ~400 ctypes launches"),
        clocks=clocks,
        model_flops=dict(per_sample_fwd_bwd_all_rows_GF=round(wl.flop_mult * fl["total"] / 1e9, 1),
                         per_sample_fwd_bwd_executed_GF=round(wl.flop_mult * fl_done["total"] / 1e9, 1),
                         achieved_TFLOPs_per_gpu_all_rows=round(value / world * wl.flop_mult * fl["total"] / 1e12, 1),
                         achieved_TFLOPs_per_gpu_executed=round(value / world * wl.flop_mult * fl_done["total"] / 1e12, 1),
                         per_sample_fwd_bwd_executed_valid_rows_GF=round(wl.flop_mult * fl_valid["total"] / 1e9, 1),
                         achieved_TFLOPs_per_gpu_executed_valid_rows=round(value / world * wl.flop_mult * fl_valid["total"] / 1e12, 1),
                         frac_of_peak_executed=round(value / world * wl.flop_mult * fl_valid["total"] / 1e12 / pk["tflops"], 4),
                         valid_rows_per_sample=(round(wl.valid_rows[0], 1) if getattr(wl, "valid_rows", None) else None),
                         note="all_rows = the reference's work (every row through every head, SURVEY 8d-D4); executed = labelled rows only in the "
                              "heads, all S rows in the encoder; executed_valid_rows = what runs: the encoder on the rows that exist (padding-free), "
                              "frac_of_peak_executed is stated against it",
                         labelled_rows_per_sample=dict(mlm=round(labelled[0], 1), token=round(labelled[1], 1))),
    )
    line["optimizer"] = dict(kernel="adamw_kernel (multi-tensor, 1 launch)", ms_per_step=round(opt_ms, 3), elements=opt_elems,
                             achieved_GBs=round(30.0 * opt_elems / (opt_ms / 1e3) / 1e9, 1), peak_GBs=pk["hbm"],
                             note="not part of `value` (metric is fwd+bwd); 30 algorithmic bytes per element (read p,g,m,v; write p,m,v and the bf16 "
                                  "operand copy of p, so that no cast pass follows)")
    if roof:
        line["roofline"] = roof
    if sync_check is not None:
        line["grad_sync"] = sync_check
    if world == 1 and not args.no_gpu_eager and args.workload == "pretrain":
        try:
            line["gpu_eager_baseline"] = gpu_eager_baseline(args, device, value)
        except Exception as e:   # a comparison leg must never cost the line
            line["gpu_eager_baseline"] = dict(error="%s: %s" % (type(e).__name__, str(e)[:200]))
    if world == 1 and not args.no_cpu_baseline and args.workload in ("pretrain", "viewpoint_select", "turn_based"):
        line["cpu_baseline"] = cpu_baseline(args, steps=2)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    os.dup2(2, 1)
    _leave(world, graphed)


def _leave(world, graphed):
    """End of a multi-rank run.  destroy_process_group() on a communicator whose collectives live in a captured CUDA graph did not
    return on the 2-GPU box (the run printed its line and then sat in the destructor until the timeout): every rank drains its
    device, meets the others once more, and leaves the process without tearing NCCL down."""
    if world <= 1:
        return
    torch.cuda.synchronize()
    torch.distributed.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def grad_sync_check(wl, model, batch, device, world):
    """Correctness of the data-parallel exchange on real NCCL ranks (DDP's contract, vs/pretrain.py:96-102): on a fixed batch per
    rank with dropout off, (a) every rank ends the step with bit-identical gradients, (b) they equal the mean over ranks of the
    gradients each rank computes on its own (the same step with the reducer detached, averaged by one plain all-reduce)."""
    dist = torch.distributed
    mods = [model] + [m for m in (getattr(wl, "encoder", None), getattr(wl, "decoder", None)) if m is not None]
    was_training = [m.training for m in mods]
    for m in mods:
        m.eval()     # dropout off (the encoder's and the agent modules'): the two passes draw no masks
    try:
        def flat_grads():
            return torch.cat([p.grad.detach().float().reshape(-1) for p in wl.params if p.grad is not None])

        def one(sync):
            for p in wl.params:
                p.grad = None
            with torch.enable_grad():
                out = wl.forward(batch)
                out[0].backward()
                if sync:
                    wl.post_backward()      # the agents' parameters outside the flat buffer
            torch.cuda.synchronize()
            return flat_grads()
        synced = one(True)
        reducers = [(m, m._vt_reducer) for m in model.modules() if getattr(m, "_vt_reducer", None) is not None]
        for m, _ in reducers:
            m._vt_reducer = None
        try:
            local = one(False)
        finally:
            for m, r in reducers:
                m._vt_reducer = r
        mean = local.clone()
        dist.all_reduce(mean, op=dist.ReduceOp.SUM)
        mean /= world
        lo, hi = synced.clone(), synced.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        identical = bool(torch.equal(lo, hi))
        rel = float((synced - mean).norm() / mean.norm().clamp_min(1e-30))
        return dict(grad_sync_ok=bool(identical and rel < 1e-3), identical_across_ranks=identical, rel_l2_vs_mean_of_local=rel, elements=int(synced.numel()),
                    note="fixed batch per rank, dropout off; rel_l2 > 0 only from the fp32 summation order (split-K red.add, NCCL's reduction tree)")
    finally:
        for m, t in zip(mods, was_training):
            m.train(t)


def gpu_eager_baseline(args, device, our_value):
    """SURVEY 8d-D5's comparison bar: the reference's eager graph (here the pinned fp32 restatement in oracle/, all rows through
    the MLM head as the reference does) on the SAME B200 at the SAME batch, fp32 and under autocast(bf16); CUDA events, outside the
    timed regions of `value`.  Reported baseline, not a target."""
    from oracle import visitron_oracle as vo
    from visitron_b200 import synthetic
    out = {}
    oc = vo.OracleConfig(hidden_dropout_prob=args.dropout, attention_probs_dropout_prob=0.1 if args.dropout > 0 else 0.0)
    drop = vo.train_dropout if args.dropout > 0 else None
    for b in (args.batch, 8):
        try:
            P = vo.clone_params(vo.init_params(oc, 88, perturb=False), device=device)
            batch = {k: v.to(device) for k, v in synthetic.pretrain_batch(b, 88).items()}
            for name, ctx in (("fp32", torch.autocast("cuda", enabled=False)), ("autocast_bf16", torch.autocast("cuda", dtype=torch.bfloat16))):
                def one():
                    for v in P.values():
                        v.grad = None
                    with ctx:
                        vo.pretrain_step(P, oc, batch, drop)
                one()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n = 3
                e0.record()
                for _ in range(n):
                    one()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / n
                out[name] = dict(samples_per_s=round(b / (ms / 1e3), 2), ms_per_step=round(ms, 2), per_gpu_batch=b,
                                 ours_over_this=round(our_value / (b / (ms / 1e3)), 2))
            out["kind"] = "port (oracle/visitron_oracle.py on cuda:0; TF32 matmul %s)" % ("on" if torch.backends.cuda.matmul.allow_tf32 else "off")
            break
        except torch.cuda.OutOfMemoryError:
            out = {"note": "b=%d did not fit beside the product's buffers; retried smaller" % b}
            P = batch = None
            torch.cuda.empty_cache()
    return out


def cpu_baseline(args, steps, warmup=1):
    """The oracle port (fp32 PyTorch restatement of the reference's `cpu` mode, run_scripts/pretrain/pretrain_ndh_r2r.sh:7-9)
    on the host cores, bounded sample: b = 2 (the script's per-GPU batch), 1 warm-up + `steps` timed fwd+bwd."""
    from oracle import visitron_oracle as vo
    from visitron_b200 import synthetic
    cores = os.cpu_count() or 1
    threads = int(os.environ.get("VT_CPU_THREADS", cores))
    torch.set_num_threads(threads)
    oc = vo.OracleConfig(hidden_dropout_prob=args.dropout, attention_probs_dropout_prob=0.1 if args.dropout > 0 else 0.0)
    drop = vo.train_dropout if args.dropout > 0 else None
    b = max(1, int(getattr(args, "cpu_batch", 2)))
    if args.workload in ("viewpoint_select", "turn_based"):
        vs = args.workload == "viewpoint_select"
        steps_ep = VS_STEPS if vs else TB_STEPS
        P = vo.clone_params(vo.init_params(oc, 88, pretrain_heads=False, perturb=False))
        A, Dv, Dt = vo.init_agent_params(88, scale=0.02)
        A, D = vo.clone_params(A), vo.clone_params(Dv if vs else Dt)
        enc = synthetic.agent_batch(b, 88)
        trace = synthetic.viewpoint_trace(steps_ep, b, 88, candidates=VS_CAND, feat=VS_FEAT) if vs else synthetic.turn_trace(steps_ep, b, 88)

        def one():
            for v in list(P.values()) + list(A.values()) + list(D.values()):
                v.grad = None
            loss, _ = (vo.rollout_vs if vs else vo.rollout_tb)(P, oc, A, D, enc, trace, steps_ep)
            loss.backward()
        what = "%s workload at b=%d (S=511 encoder + 511-step LSTM + %d decoder steps; dropout off in the port)" % (args.workload, b, steps_ep)
    else:
        P = vo.clone_params(vo.init_params(oc, 88, perturb=False))
        batch = synthetic.pretrain_batch(b, 88)

        def one():
            vo.pretrain_step(P, oc, batch, drop)
        what = "pre-training workload at b=%d (S=767, all rows through the MLM head as the reference does)" % b
    for _ in range(max(1, warmup)):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt = (time.perf_counter() - t0) / steps
    return dict(value=round(b / dt, 4), unit=UNIT, cores=threads, kind="port",
                sample="%d fwd+bwd steps of the same %s, fp32 torch CPU, %d threads of %d host cores; %.2f s/step" % (steps, what, threads, cores, dt))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # every step is a bounded sample (b = 2) of the workload, ~2.5 s on 16 host cores: --steps / --warmup are honoured up to 40 / 5
    # so that the whole run stays within a couple of minutes
    steps, warm = max(1, min(args.steps, 40)), max(1, min(args.warmup, 5))
    cb = cpu_baseline(args, steps, warm)
    line = dict(impl="reference", metric=METRIC, value=cb["value"], unit=UNIT, n_gpus=args.gpus, steps=steps, warmup=warm,
                ms_per_step=round(args.cpu_batch / cb["value"] * 1e3, 1), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=make_config(args, args.gpus),
                cpu_baseline=cb, e2e=dict(value=cb["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="per-GPU batch (the reference scripts use 2; 32 fills the tensor cores)")
    ap.add_argument("--dropout", type=float, default=0.3, help="hidden dropout (run_scripts/pretrain/pretrain_ndh_r2r.sh:43 uses 0.3)")
    ap.add_argument("--workload", default="pretrain", choices=["pretrain", "viewpoint_select", "turn_based", "qa_classifier"],
                    help="pretrain = BASELINE configs[2] (the data-parallel headline); viewpoint_select = configs[1] (encoder at S=511 + candidate scorer); "
                         "turn_based = configs[3]; qa_classifier = configs[4] (sweep --batch)")
    ap.add_argument("--min-warm-s", type=float, default=1.0, help="keep warming up until this much wall time has passed")
    ap.add_argument("--repeats", type=int, default=5, help="timed regions of --steps steps each; the median is reported, all are listed")
    ap.add_argument("--manual-gc", action="store_true", help="diagnosis: gc.freeze() + gc.disable() around the timed regions")
    ap.add_argument("--feat-bf16", action="store_true", help="region features / location embeddings come from a bf16 cache (pretrain workload)")
    ap.add_argument("--no-clocks", action="store_true", help="diagnosis: do not sample clocks during the timed region")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-batch", type=int, default=2, help="batch of the CPU baseline's bounded sample (the scripts' per-GPU batch is 2; BASELINE C1 also asks for 8)")
    ap.add_argument("--no-gpu-eager", action="store_true", help="skip the oracle-on-the-GPU comparison leg (SURVEY 8d-D5)")
    ap.add_argument("--dense", action="store_true", help="pre-training on the dense [B, S] layout (padding rows included) instead of packed rows")
    ap.add_argument("--no-graph", action="store_true", help="time the eager path (~400 ctypes launches per step) instead of the captured CUDA graph")
    ap.add_argument("--no-roofline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
