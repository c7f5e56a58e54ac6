"""Import shim for the UNMODIFIED reference modules (build container only)  --  TEST INFRASTRUCTURE.

/root/reference is a read-only mount that exists in the build container and NOT on the GPU box,
so nothing that runs there (``-m gpu`` tests, smoke(), bench.py) may call this.  It is used by
``oracle/make_golden.py`` (fixture generation) and by the ``not gpu`` tests that pin
``oracle/visitron_oracle.py`` against the real reference (skipped when the mount is absent).

What has to be worked around:
  * site-packages holds a modern ``transformers`` that shadows the reference's vendored
    ``transformers/pytorch_transformers`` namespace dir -> pre-seed ``sys.modules['transformers']``;
  * ``hf/file_utils.py:19-21`` imports boto3/botocore (absent) -> stub modules;
  * reference modules import one another as top-level names (``encoder``, ``agent_models``) ->
    put the task directory on ``sys.path``;
  * ``params.py`` parses ``sys.argv`` at import and the data loaders need MatterSim/lmdb -> never imported.
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("VISITRON_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "tasks", "viewpoint_select"))


_loaded = {}


def load():
    """Returns a namespace with the reference classes on the hot path."""
    if _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError("reference tree not mounted at %s" % REFERENCE_ROOT)
    saved_tf = {k: v for k, v in sys.modules.items() if k == "transformers" or k.startswith("transformers.")}
    for k in saved_tf:
        del sys.modules[k]
    pkg = types.ModuleType("transformers")
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "transformers")]
    sys.modules["transformers"] = pkg
    for n in ("boto3", "botocore", "botocore.exceptions"):
        sys.modules.setdefault(n, types.ModuleType(n))
    sys.modules["botocore.exceptions"].ClientError = Exception
    sys.modules["botocore"].exceptions = sys.modules["botocore.exceptions"]
    vs = os.path.join(REFERENCE_ROOT, "tasks", "viewpoint_select")
    tb = os.path.join(REFERENCE_ROOT, "tasks", "turn_based")
    sys.path[:0] = [REFERENCE_ROOT, vs]
    try:
        hf = importlib.import_module("transformers.pytorch_transformers.modeling_bert")
        opt = importlib.import_module("transformers.pytorch_transformers.optimization")
        enc = importlib.import_module("encoder")            # tasks/viewpoint_select/encoder.py
        vs_models = importlib.import_module("agent_models")  # tasks/viewpoint_select/agent_models.py
        spec = importlib.util.spec_from_file_location("tb_agent_models", os.path.join(tb, "agent_models.py"))
        tb_models = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(tb_models)
    finally:
        sys.path.remove(REFERENCE_ROOT)
        sys.path.remove(vs)
    ns = types.SimpleNamespace(
        BertConfig=hf.BertConfig, hf=hf, optimization=opt, encoder=enc, vs_models=vs_models, tb_models=tb_models,
        PreTrainOscar=enc.PreTrainOscar, BertImgModelwithLocationEmbeds=enc.BertImgModelwithLocationEmbeds)
    _loaded["ns"] = ns
    # keep the vendored namespace registered under a private alias and restore the real package
    for k in [k for k in sys.modules if k == "transformers" or k.startswith("transformers.")]:
        sys.modules["_visitron_ref_" + k] = sys.modules.pop(k)
    sys.modules.update(saved_tf)
    return ns


def make_config(ns, cfg):
    """BertConfig patched the way vs/model_utils.py:75-83 does, from an OracleConfig."""
    c = ns.BertConfig(cfg.vocab_size)
    for k in ("hidden_size", "num_hidden_layers", "num_attention_heads", "intermediate_size",
              "max_position_embeddings", "type_vocab_size", "layer_norm_eps", "initializer_range",
              "img_feature_dim", "use_img_layernorm", "img_layer_norm_eps", "action_space",
              "detector_classes", "hidden_dropout_prob", "attention_probs_dropout_prob"):
        setattr(c, k, getattr(cfg, k))
    c.img_feature_type = "faster_r-cnn"
    c.hidden_act = "gelu"
    return c


def build_pretrain(ns, cfg, params):
    """Reference PreTrainOscar carrying exactly ``params`` (a state_dict-keyed dict)."""
    m = ns.PreTrainOscar(make_config(ns, cfg))
    sd = {k: v.detach().clone().float() for k, v in params.items()}
    sd["mlmhead.predictions.decoder.weight"] = sd["bert.embeddings.word_embeddings.weight"]
    missing, unexpected = m.load_state_dict(sd, strict=True), None
    m.tie_weights()
    return m


def load_mask_tokens():
    """The reference's ``PretrainDataset._mask_tokens`` (tasks/viewpoint_select/data_loader_pretrain.py:549-612) as an unbound
    function; the data-loader module's unrelated imports that are missing here (lmdb, networkx, ...) are stubbed."""
    load()
    vs = os.path.join(REFERENCE_ROOT, "tasks", "viewpoint_select")
    for n in ("lmdb", "networkx", "jsonlines", "tensorboardX", "MatterSim"):
        try:
            importlib.import_module(n)
        except Exception:
            sys.modules.setdefault(n, types.ModuleType(n))
    sys.path[:0] = [vs]
    try:
        mod = importlib.import_module("data_loader_pretrain")
    finally:
        sys.path.remove(vs)
    return mod.PretrainDataset._mask_tokens


def load_preprocess_item():
    """``PretrainDataset._preprocess_item`` (data_loader_pretrain.py:636-712), unbound."""
    load_mask_tokens()
    return sys.modules["data_loader_pretrain"].PretrainDataset._preprocess_item
