import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from visitron_b200 import modeling, synthetic
from oracle import visitron_oracle as vo
from parity_common import load_params, make_config
DEV = "cuda"
oc = vo.OracleConfig(num_hidden_layers=2)
model = modeling.PreTrainOscar(make_config(oc)); load_params(model, vo.init_params(oc, 9)); model.to(DEV).train()
batch = {k: v.to(DEV) for k, v in synthetic.pretrain_batch(8, 88).items()}
res = {}
for packed in (False, True, False):
    model.vt_packed = packed
    for p in model.parameters(): p.grad = None
    out = model(**batch); out[0].backward(); torch.cuda.synchronize()
    res.setdefault(packed, []).append((torch.stack([o.detach() for o in out]).double().cpu(), {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}))
    del out
(o0, g0), (o0b, g0b) = res[False]
(o1, g1), = res[True]
print("outputs dense ", o0.tolist()); print("outputs packed", o1.tolist()); print("dense run-to-run", (o0 - o0b).abs().max().item())
rows = []
for k in g0:
    d = float((g1[k] - g0[k]).norm()) / max(float(g0[k].norm()), 1e-12)
    dd = float((g0b[k] - g0[k]).norm()) / max(float(g0[k].norm()), 1e-12)
    rows.append((d, dd, k))
for d, dd, k in sorted(rows, reverse=True)[:14]: print("%.3e (dense run-to-run %.1e)  %s" % (d, dd, k))
