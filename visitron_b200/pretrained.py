"""Configuration object and checkpoint plumbing the reference's loaders expect from a model class.

``load_oscar_weights`` (tasks/viewpoint_select/model_utils.py:36-111) builds a ``BertConfig``, patches
attributes onto it, then calls ``ModelClass.from_pretrained(path, config=config)`` and later
``save_pretrained``; ``train.py`` takes ``model.bert``.  This module provides that contract for the
B200 classes without importing the reference: a config that accepts arbitrary attributes and
round-trips through ``config.json``, and a base class with ``from_pretrained`` / ``save_pretrained``
/ ``init_weights`` / weight tying / embedding resizing that reads and writes the same
``pytorch_model.bin`` state_dict files (hf/modeling_utils.py:219-520 documents the behaviour kept:
gamma/beta key renaming, ``bert.`` prefix juggling, non-strict load, eval() after load).
Any object with the same attributes (e.g. the reference's own ``BertConfig``) is accepted as config.
"""
from __future__ import annotations

import copy
import json
import logging
import os

import torch
from torch import nn

logger = logging.getLogger(__name__)

CONFIG_NAME = "config.json"
WEIGHTS_NAME = "pytorch_model.bin"


class BertConfig:
    """bert-base defaults plus the VISITRON additions (vs/model_utils.py:75-83)."""

    def __init__(self, vocab_size_or_config_json_file=30522, hidden_size=768, num_hidden_layers=12,
                 num_attention_heads=12, intermediate_size=3072, hidden_act="gelu", hidden_dropout_prob=0.1,
                 attention_probs_dropout_prob=0.1, max_position_embeddings=512, type_vocab_size=2,
                 initializer_range=0.02, layer_norm_eps=1e-12, **kwargs):
        self.output_attentions = False
        self.output_hidden_states = False
        self.torchscript = False
        self.img_feature_dim = 2054
        self.img_feature_type = "faster_r-cnn"
        self.use_img_layernorm = 0
        self.img_layer_norm_eps = 1e-12
        self.action_space = 36
        self.detector_classes = 1601
        if isinstance(vocab_size_or_config_json_file, str):
            with open(vocab_size_or_config_json_file, "r", encoding="utf-8") as f:
                for k, v in json.load(f).items():
                    setattr(self, k, v)
        else:
            self.vocab_size = vocab_size_or_config_json_file
            self.hidden_size = hidden_size
            self.num_hidden_layers = num_hidden_layers
            self.num_attention_heads = num_attention_heads
            self.intermediate_size = intermediate_size
            self.hidden_act = hidden_act
            self.hidden_dropout_prob = hidden_dropout_prob
            self.attention_probs_dropout_prob = attention_probs_dropout_prob
            self.max_position_embeddings = max_position_embeddings
            self.type_vocab_size = type_vocab_size
            self.initializer_range = initializer_range
            self.layer_norm_eps = layer_norm_eps
        for k, v in kwargs.items():
            setattr(self, k, v)

    @classmethod
    def from_pretrained(cls, path, **kwargs):
        f = os.path.join(path, CONFIG_NAME) if os.path.isdir(path) else path
        cfg = cls(f)
        for k, v in kwargs.items():
            setattr(cfg, k, v)
        return cfg

    def to_dict(self):
        return copy.deepcopy(self.__dict__)

    def to_json_string(self):
        return json.dumps(self.to_dict(), indent=2, sort_keys=True) + "\n"

    def save_pretrained(self, save_directory):
        with open(os.path.join(save_directory, CONFIG_NAME), "w", encoding="utf-8") as f:
            f.write(self.to_json_string())

    def __repr__(self):
        return "BertConfig " + self.to_json_string()


def check_envelope(config):
    """The CUDA path is specialised for bert-base geometry; anything else is refused, not emulated."""
    if config.hidden_size % config.num_attention_heads != 0:
        # same error the reference raises (hf/modeling_bert.py:277-280)
        raise ValueError("The hidden size (%d) is not a multiple of the number of attention heads (%d)"
                         % (config.hidden_size, config.num_attention_heads))
    problems = []
    if config.hidden_size != 768:
        problems.append("hidden_size=%d (768)" % config.hidden_size)
    if config.hidden_size // config.num_attention_heads != 64:
        problems.append("head dim %d (64)" % (config.hidden_size // config.num_attention_heads))
    if config.intermediate_size % 64 != 0:
        problems.append("intermediate_size=%d (multiple of 64)" % config.intermediate_size)
    act = getattr(config, "hidden_act", "gelu")
    if act != "gelu":
        problems.append("hidden_act=%r ('gelu')" % (act,))
    if getattr(config, "output_attentions", False):
        problems.append("output_attentions=True (would materialise the S x S probabilities)")
    if problems:
        raise NotImplementedError("visitron_b200 supports the bert-base envelope only; got " + ", ".join(problems))


class PreTrainedModel(nn.Module):
    config_class = BertConfig
    base_model_prefix = "bert"

    def __init__(self, config, *inputs, **kwargs):
        super().__init__()
        for attr in ("hidden_size", "num_attention_heads", "num_hidden_layers", "vocab_size"):
            if not hasattr(config, attr):
                raise ValueError("Parameter config in `%s(config)` must carry `%s`" % (type(self).__name__, attr))
        self.config = config

    # -- initialisation (hf/modeling_bert.py:550-561) ---------------------------------------
    def init_weights(self, module):
        from .modeling import BertLayerNorm
        if isinstance(module, (nn.Linear, nn.Embedding)):
            module.weight.data.normal_(mean=0.0, std=self.config.initializer_range)
        elif isinstance(module, BertLayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)
        if isinstance(module, nn.Linear) and module.bias is not None:
            module.bias.data.zero_()

    def _tie_or_clone_weights(self, first_module, second_module):
        if getattr(self.config, "torchscript", False):
            first_module.weight = nn.Parameter(second_module.weight.clone())
        else:
            first_module.weight = second_module.weight

    def _get_resized_embeddings(self, old_embeddings, new_num_tokens=None):
        if new_num_tokens is None:
            return old_embeddings
        old_n, dim = old_embeddings.weight.size()
        if old_n == new_num_tokens:
            return old_embeddings
        new = nn.Embedding(new_num_tokens, dim)
        new.to(old_embeddings.weight.device)
        self.init_weights(new)
        n = min(old_n, new_num_tokens)
        new.weight.data[:n, :] = old_embeddings.weight.data[:n, :]
        return new

    def prune_heads(self, heads_to_prune):
        raise NotImplementedError("head pruning changes the 12 x 64 head geometry the CUDA kernels are specialised for")

    # -- checkpoints ------------------------------------------------------------------------
    def save_pretrained(self, save_directory):
        assert os.path.isdir(save_directory), "Saving path should be a directory where the model and configuration can be saved"
        model_to_save = self.module if hasattr(self, "module") else self
        cfg = model_to_save.config
        if hasattr(cfg, "save_pretrained"):
            cfg.save_pretrained(save_directory)
        else:  # the reference's own config object
            with open(os.path.join(save_directory, CONFIG_NAME), "w", encoding="utf-8") as f:
                f.write(cfg.to_json_string())
        torch.save(model_to_save.state_dict(), os.path.join(save_directory, WEIGHTS_NAME))

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path, *model_args, **kwargs):
        config = kwargs.pop("config", None)
        state_dict = kwargs.pop("state_dict", None)
        output_loading_info = kwargs.pop("output_loading_info", False)
        kwargs.pop("cache_dir", None)
        kwargs.pop("from_tf", None)
        if config is None:
            config = cls.config_class.from_pretrained(pretrained_model_name_or_path, **kwargs)
        model = cls(config, *model_args)
        if state_dict is None:
            path = pretrained_model_name_or_path
            if os.path.isdir(path):
                path = os.path.join(path, WEIGHTS_NAME)
            if not os.path.isfile(path):
                raise EnvironmentError("no checkpoint at %s (model names cannot be downloaded here)" % path)
            state_dict = torch.load(path, map_location="cpu")
        # TF-era names
        renamed = {}
        for k in list(state_dict.keys()):
            nk = k.replace("gamma", "weight") if "gamma" in k else k
            nk = nk.replace("beta", "bias") if "beta" in nk else nk
            renamed[nk] = state_dict[k]
        state_dict = renamed
        prefix = cls.base_model_prefix
        has_prefix = any(s.startswith(prefix + ".") for s in state_dict)
        target = model
        if not hasattr(model, prefix) and has_prefix:          # loading a head checkpoint into the bare encoder
            state_dict = {k[len(prefix) + 1:]: v for k, v in state_dict.items() if k.startswith(prefix + ".")}
        elif hasattr(model, prefix) and not has_prefix:        # loading a bare encoder into a model with heads
            target = getattr(model, prefix)
        result = target.load_state_dict(state_dict, strict=False)
        missing, unexpected = list(result.missing_keys), list(result.unexpected_keys)
        if missing:
            logger.info("Weights of %s not initialized from pretrained model: %s", cls.__name__, missing)
        if unexpected:
            logger.info("Weights from pretrained model not used in %s: %s", cls.__name__, unexpected)
        if hasattr(model, "tie_weights"):
            model.tie_weights()
        model.eval()
        if output_loading_info:
            return model, {"missing_keys": missing, "unexpected_keys": unexpected, "error_msgs": []}
        return model
