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
import subprocess
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
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    power_w_max=max(pw) if pw else None, samples=len(sm), reasons=sorted(reasons))


def build_model(args, device):
    from visitron_b200 import modeling
    from visitron_b200.pretrained import BertConfig
    cfg = BertConfig(V, hidden_dropout_prob=args.dropout, attention_probs_dropout_prob=0.1 if args.dropout > 0 else 0.0)
    cfg.img_feature_dim, cfg.detector_classes, cfg.action_space = IMG, DC, A
    torch.manual_seed(88)
    if args.workload == "viewpoint_select":
        model = modeling.BertImgModelwithLocationEmbeds(cfg)   # what train.py hands the agent as `model.bert` (vs/train.py:47)
    else:
        model = modeling.PreTrainOscar(cfg)     # random-init bert-base (hf/modeling_bert.py:550-561), no checkpoint
    return model.to(device).train()


VS_STEPS, VS_CAND, VS_FEAT, VS_HID = 10, 12, 2052, 512   # decoder steps, candidates incl. STOP, 2048+4 features, LSTM width


class Workload:
    """One step of the hot path for `--workload`: host batches, the step function and the algorithmic FLOPs."""

    def __init__(self, args, device, rank, world):
        from visitron_b200 import agent_heads, synthetic
        self.args, self.device, self.world = args, device, world
        self.model = build_model(args, device)
        b, n_host = args.batch, 4
        self.extra_params = []
        if args.workload == "viewpoint_select":
            # BASELINE configs[1]: the encoder call of OscarEncoder (A13, S = 511 text-only, token types, the ~uint8 mask)
            # + the candidate scorer with its masked CE (A14) at each of the 10 decoder steps.  The stock torch.nn LSTM
            # encoder / decoder between them is SURVEY 8f-N1 (not built): the decoder state of step t is stood in for by
            # sequence_output[:, t, :512], which keeps every head gradient flowing into the encoder.
            g = torch.Generator().manual_seed(88)
            self.w_in = torch.nn.Parameter((0.02 * torch.randn(VS_FEAT, VS_HID, generator=g)).to(device))
            self.extra_params = [self.w_in]
            host = []
            for i in range(n_host):
                hb = synthetic.finetune_encoder_batch(b, seed=88 + rank + 1000 * i)
                steps = [synthetic.viewpoint_step_batch(b, seed=88 + rank + 1000 * i + 17 * t, candidates=VS_CAND, feat=VS_FEAT) for t in range(VS_STEPS)]
                hb["cand_feat"] = torch.stack([s["cand_feat"] for s in steps])       # [10, b, 12, 2052]
                hb["cand_len"] = torch.stack([s["cand_len"] for s in steps])
                hb["target"] = torch.stack([s["target"] for s in steps])
                host.append(hb)
            self.S, self.R = 511, 0
            self.labelled = (0.0, 0.0)
            self.heads_flops = VS_STEPS * (2 * VS_HID * VS_FEAT + 2 * VS_CAND * VS_FEAT)
            self.desc = ("viewpoint_select_step: BertImgModelwithLocationEmbeds fwd+bwd as OscarEncoder calls it (T=511 text-only, token types, "
                         "~uint8 mask) + SoftDotAttention candidate scorer + masked CE at 10 decoder steps (12 candidates x 2052), bert-base "
                         "random init [BASELINE configs[1]; the stock LSTM encoder/decoder of the agent is not part of the path, SURVEY 8f-N1]")
        else:
            host = [synthetic.pretrain_batch(b, seed=88 + rank + 1000 * i) for i in range(n_host)]  # per-rank shard, seed 88+rank
            self.S, self.R = 767, 256
            self.labelled = (sum(int((hb["labels"] != -1).sum()) for hb in host) / (n_host * b),
                             sum(int((hb["token_labels"] != -1).sum()) for hb in host) / (n_host * b))
            self.heads_flops = None
            self.desc = ("pretrain_step: PreTrainOscar fwd+bwd, bert-base random init, NDH-shaped batch (T=511 text + R=256 region slots, "
                         "img_feature_dim 2054, 15% MLM + masked detector-class + 36-way action labels), all objectives [BASELINE configs[2]; "
                         "configs[1] = --workload viewpoint_select]")
        self.host = [{k: v.pin_memory() for k, v in hb.items()} for hb in host]
        self.n_host = n_host
        self.params = list(self.model.parameters()) + self.extra_params
        self._heads = agent_heads

    def step(self, batch):
        for p in self.params:
            p.grad = None
        if self.args.workload != "viewpoint_select":
            out = self.model(**batch)
            out[0].backward()
            return out
        seq, _ = self.model(batch["input_ids"], token_type_ids=batch["token_type_ids"], attention_mask=batch["attention_mask"])
        loss = None
        for t in range(VS_STEPS):
            _, l, _ = self._heads.candidate_scores_ce(self.w_in, seq[:, t, :VS_HID], batch["cand_feat"][t], batch["cand_len"][t], batch["target"][t])
            loss = l if loss is None else loss + l
        loss.backward()
        if self.world > 1:   # the scorer's weight is outside the encoder's flat buffer: one small all-reduce of its own
            torch.distributed.all_reduce(self.w_in.grad, op=torch.distributed.ReduceOp.AVG)
        return (loss,)

    def flops(self, executed=False):
        fl = flops_per_sample(self.S, self.R, all_rows_heads=not executed, labelled=self.labelled)
        if self.heads_flops is not None:   # no pre-training heads, no image projection on this workload
            fl["total"] += self.heads_flops - fl["heads"] - fl["img"]
            fl["heads"], fl["img"] = self.heads_flops, 0
        return fl


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

    for i in range(max(args.warmup, 3)):
        step(dev_batches[i % n_host])
    # ---- device-resident throughput (`value`) ----
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    lib.reset_launch_count()
    ms = timed(lambda i: step(dev_batches[i % n_host]), args.steps)
    launches = lib.launch_count()
    clocks = sampler.stop() if sampler else None
    value = world * b * args.steps / (ms / 1e3)

    # ---- end to end from pinned host memory through the public API (`e2e`) ----
    copy_stream = torch.cuda.Stream()
    h2d_bytes = sum(v.numel() * v.element_size() for v in host[0].values())
    slots = [None, None]

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            slots[i % 2] = ({k: v.to(device, non_blocking=True) for k, v in host[i % n_host].items()}, copy_stream.record_event())

    def e2e_step(i):
        if slots[i % 2] is None:
            prefetch(i)
        batch, ev = slots[i % 2]
        slots[i % 2] = None
        prefetch(i + 1)                       # next batch's H2D overlaps this step's compute
        torch.cuda.current_stream().wait_event(ev)
        out = step(batch)
        for t in batch.values():
            t.record_stream(torch.cuda.current_stream())
        return torch.stack([o.detach().float() for o in out]).cpu()   # D2H read of the step's scalars (host sync, as the loop's .item())

    e2e_step(0)
    slots = [None, None]
    e2e_ms = timed(e2e_step, args.steps)
    e2e_value = world * b * args.steps / (e2e_ms / 1e3)

    # ---- roofline of the dominant kernel: per-launch CUDA events around every GEMM, same K steps ----
    roof = None
    if not args.no_roofline:
        lib.GEMM_EVENTS = []
        timed(lambda i: step(dev_batches[i % n_host]), args.steps)
        evs, lib.GEMM_EVENTS = lib.GEMM_EVENTS, None
        torch.cuda.synchronize()
        tot_ms = sum(a.elapsed_time(c) for a, c, _ in evs)
        tot_fl = sum(f for _, _, f in evs)
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
                    note="algorithmic 2*M*N*K per launch / CUDA-event time per launch, separate pass of the same K steps")

    # ---- the step after backward, reported separately (the metric is fwd+bwd): fused AdamW, one launch ----
    from visitron_b200 import optimization as vopt
    no_decay = ["bias", "LayerNorm.weight"]
    named = [(n, p) for n, p in model.named_parameters() if p.grad is not None]
    opt = vopt.AdamW([dict(params=[p for n, p in named if not any(nd in n for nd in no_decay)], weight_decay=0.01),
                      dict(params=[p for n, p in named if any(nd in n for nd in no_decay)], weight_decay=0.0)], lr=5e-5, eps=1e-8)
    opt.step()
    opt_ms = timed(lambda i: opt.step(), 5) / 5
    opt_elems = sum(p.numel() for _, p in named)

    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return
    S = wl.S
    fl, fl_done = wl.flops(), wl.flops(executed=True)
    n_out = 7 if args.workload == "pretrain" else 1
    pk = peaks()
    line = dict(
        metric=METRIC, value=round(value, 2), unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
        ms_per_step=round(ms / args.steps, 3), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16",
        data="synthetic",
        config=dict(workload=wl.desc,
                    per_gpu_batch=b, global_batch=b * world, seq_len=S, parallelism="dp%d" % world, dropout=args.dropout,
                    l2="per-step working set (%.1f GB of activations) >> 126 MB L2; 4 rotating batches" % (b * S * 24.5e3 * 12 / 1e9),
                    grad_allreduce="fp32, one bucket per encoder layer, overlapped with backward" if world > 1 else "none (1 GPU)"),
        e2e=dict(value=round(e2e_value, 2), unit=UNIT, h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=4 * n_out,
                 ms_per_step=round(e2e_ms / args.steps, 3)),
        gpu_launches=int(launches),
        clocks=clocks,
        model_flops=dict(per_sample_fwd_bwd_all_rows_GF=round(3 * fl["total"] / 1e9, 1),
                         per_sample_fwd_bwd_executed_GF=round(3 * fl_done["total"] / 1e9, 1),
                         achieved_TFLOPs_per_gpu_all_rows=round(value / world * 3 * fl["total"] / 1e12, 1),
                         achieved_TFLOPs_per_gpu_executed=round(value / world * 3 * fl_done["total"] / 1e12, 1),
                         frac_of_peak_executed=round(value / world * 3 * fl_done["total"] / 1e12 / pk["tflops"], 4),
                         labelled_rows_per_sample=dict(mlm=round(labelled[0], 1), token=round(labelled[1], 1))),
    )
    line["optimizer"] = dict(kernel="adamw_kernel (multi-tensor, 1 launch)", ms_per_step=round(opt_ms, 3), elements=opt_elems,
                             achieved_GBs=round(28.0 * opt_elems / (opt_ms / 1e3) / 1e9, 1), peak_GBs=pk["hbm"],
                             note="not part of `value` (metric is fwd+bwd); 28 algorithmic bytes per element (read p,g,m,v; write p,m,v)")
    if roof:
        line["roofline"] = roof
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args, steps=2)
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


def cpu_baseline(args, steps):
    """The oracle port (fp32 PyTorch restatement of the reference's `cpu` mode, run_scripts/pretrain/pretrain_ndh_r2r.sh:7-9)
    on the host cores, bounded sample: b = 2 (the script's per-GPU batch), 1 warm-up + `steps` timed fwd+bwd."""
    from oracle import visitron_oracle as vo
    from visitron_b200 import synthetic
    cores = os.cpu_count() or 1
    threads = int(os.environ.get("VT_CPU_THREADS", cores))
    torch.set_num_threads(threads)
    oc = vo.OracleConfig(hidden_dropout_prob=args.dropout, attention_probs_dropout_prob=0.1 if args.dropout > 0 else 0.0)
    drop = vo.train_dropout if args.dropout > 0 else None
    b = 2
    if args.workload == "viewpoint_select":
        P = vo.clone_params(vo.init_params(oc, 88, pretrain_heads=False, perturb=False))
        w_in = vo.clone_params(vo.init_head_params(88))["candidate_att_layer.linear_in.weight"]
        enc = synthetic.finetune_encoder_batch(b, 88)
        vsb = [synthetic.viewpoint_step_batch(b, 88 + 17 * t, candidates=VS_CAND, feat=VS_FEAT) for t in range(VS_STEPS)]

        def one():
            for v in list(P.values()) + [w_in]:
                v.grad = None
            seq, _ = vo.bert_img_forward(P, oc, enc["input_ids"], enc["token_type_ids"], enc["attention_mask"], drop=drop)
            loss = sum(vo.candidate_scores_ce(w_in, seq[:, t, :VS_HID], s["cand_feat"], s["cand_len"], s["target"])[1] for t, s in enumerate(vsb))
            loss.backward()
        what = "viewpoint_select workload at b=%d (S=511 encoder + 10 candidate-scorer steps)" % b
    else:
        P = vo.clone_params(vo.init_params(oc, 88, perturb=False))
        batch = synthetic.pretrain_batch(b, 88)

        def one():
            vo.pretrain_step(P, oc, batch, drop)
        what = "pre-training workload at b=%d (S=767, all rows through the MLM head as the reference does)" % b
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
    steps = max(1, min(args.steps, 5))
    cb = cpu_baseline(args, steps)
    line = dict(impl="reference", metric=METRIC, value=cb["value"], unit=UNIT, n_gpus=args.gpus, steps=steps, warmup=1,
                ms_per_step=round(2 / cb["value"] * 1e3, 1), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload="%s step (same workload as the default arm; bounded sample b=2 per step on the host CPU)" % args.workload,
                            per_gpu_batch=2, seq_len=767 if args.workload == "pretrain" else 511),
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
    ap.add_argument("--workload", default="pretrain", choices=["pretrain", "viewpoint_select"],
                    help="pretrain = BASELINE configs[2] (the data-parallel headline); viewpoint_select = configs[1] (encoder at S=511 + candidate scorer)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
