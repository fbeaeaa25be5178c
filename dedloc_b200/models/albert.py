"""ALBERT for pre-training (MLM + sentence-order prediction) with Hugging Face-compatible parameter names.

The reference builds `transformers.AlbertForPreTraining(config)` (albert/run_trainer.py:56-70) and saves /
loads it with `save_pretrained` / `from_pretrained` (run_first_peer.py:131-134). This module re-implements the
same network B200-first:
  * the module tree reproduces the HF names (`albert.encoder.albert_layer_groups.0.albert_layers.0.*`,
    tied `predictions.decoder`), so `state_dict()` round-trips through HF checkpoints;
  * one set of layer weights is applied `num_hidden_layers` times (weight sharing);
  * linears run on the tcgen05 GEMM with bias / gelu_new / residual fused in the epilogue and weight gradients
    accumulated over the 24 uses straight into fp32 main-grad buffers (ops/gemm.py), when `impl="native"`;
  * the MLM head gathers the masked positions BEFORE the vocabulary GEMM instead of computing
    logits for all 512 positions.
`impl="torch"` is the plain PyTorch path (CPU tests, oracle).
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import asdict, dataclass, field
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import cross_entropy as _ce
from ..ops.elementwise import gelu_new


@dataclass
class AlbertConfig:
    vocab_size: int = 30000
    embedding_size: int = 128
    hidden_size: int = 1024
    num_hidden_layers: int = 24
    num_hidden_groups: int = 1
    num_attention_heads: int = 16
    intermediate_size: int = 4096
    inner_group_num: int = 1
    hidden_act: str = "gelu_new"
    hidden_dropout_prob: float = 0.0
    attention_probs_dropout_prob: float = 0.0
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    initializer_range: float = 0.02
    layer_norm_eps: float = 1e-12
    classifier_dropout_prob: float = 0.1
    pad_token_id: int = 0
    bos_token_id: int = 2
    eos_token_id: int = 3
    model_type: str = "albert"
    architectures: list = field(default_factory=lambda: ["AlbertForPreTraining"])

    @classmethod
    def albert_large_v2(cls, **kw) -> "AlbertConfig":
        """albert-large-v2 dimensions (the config URL at albert/arguments.py:97-99)."""
        return cls(**kw)

    @classmethod
    def tiny(cls, **kw) -> "AlbertConfig":
        d = dict(vocab_size=1024, embedding_size=32, hidden_size=64, num_hidden_layers=2, num_attention_heads=4,
                 intermediate_size=128, max_position_embeddings=128)
        d.update(kw)
        return cls(**d)

    @classmethod
    def small(cls, **kw) -> "AlbertConfig":
        """Small but tcgen05-friendly dimensions (every width a multiple of 64) for GPU smoke tests."""
        d = dict(vocab_size=2048, embedding_size=128, hidden_size=256, num_hidden_layers=4, num_attention_heads=4,
                 intermediate_size=1024, max_position_embeddings=128)
        d.update(kw)
        return cls(**d)

    @classmethod
    def from_dict(cls, d: Dict) -> "AlbertConfig":
        known = {f for f in cls.__dataclass_fields__}
        return cls(**{k: v for k, v in d.items() if k in known})

    @classmethod
    def from_pretrained(cls, path: str, **_) -> "AlbertConfig":
        if os.path.isdir(path):
            path = os.path.join(path, "config.json")
        with open(path) as f:
            return cls.from_dict(json.load(f))

    def to_dict(self) -> Dict:
        return asdict(self)

    def save_pretrained(self, directory: str) -> None:
        os.makedirs(directory, exist_ok=True)
        with open(os.path.join(directory, "config.json"), "w") as f:
            json.dump(self.to_dict(), f, indent=2, sort_keys=True)


class _Linear(nn.Module):
    """nn.Linear-shaped parameters (`weight` [out,in], `bias` [out]); the compute path is chosen by the owner."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features))


def _main_grad(p: torch.Tensor) -> Optional[torch.Tensor]:
    return getattr(p, "main_grad", None)


def _native_ok(x, lin, impl: str) -> bool:
    """The tcgen05 GEMM needs every reduction dimension (in_features for forward, out_features for dgrad, tokens for
    wgrad) to be a multiple of its 64-element K block."""
    return (impl == "native" and x.is_cuda and x.dtype == torch.bfloat16 and _main_grad(lin.weight) is not None
            and lin.in_features % 64 == 0 and lin.out_features % 64 == 0
            and (x.numel() // lin.in_features) % 64 == 0)


def fused_linear(x, lin: _Linear, impl: str, act: str = "none", residual=None):
    """y = act(x W^T + b) (+ residual) through the tcgen05 GEMM when impl == 'native' and the parameters carry
    fp32 `main_grad` buffers; otherwise plain PyTorch."""
    if _native_ok(x, lin, impl):
        from ..ops.gemm import linear as native_linear

        return native_linear(x, lin.weight, lin.bias, act=act, residual=residual,
                             wgrad_out=lin.weight.main_grad, bgrad_out=lin.bias.main_grad)
    y = F.linear(x, lin.weight, lin.bias)
    if act == "gelu":
        y = gelu_new(y.float()).to(y.dtype) if y.dtype != torch.float32 else gelu_new(y)
    if residual is not None:
        y = y + residual
    return y


def _ln(x, ln: nn.LayerNorm, impl: str):
    if impl == "native" and x.is_cuda:
        from ..ops.layernorm import layer_norm

        return layer_norm(x, ln, impl)
    return ln(x)


class AlbertEmbeddings(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.word_embeddings = nn.Embedding(cfg.vocab_size, cfg.embedding_size, padding_idx=cfg.pad_token_id)
        self.position_embeddings = nn.Embedding(cfg.max_position_embeddings, cfg.embedding_size)
        self.token_type_embeddings = nn.Embedding(cfg.type_vocab_size, cfg.embedding_size)
        self.LayerNorm = nn.LayerNorm(cfg.embedding_size, eps=cfg.layer_norm_eps)
        self.dropout_p = cfg.hidden_dropout_prob

    def forward(self, input_ids, token_type_ids, impl: str = "torch"):
        S = input_ids.shape[1]
        if impl == "native" and input_ids.is_cuda:
            from ..ops import embedding as _emb

            if _emb.embed_supported(self, input_ids) and S <= self.position_embeddings.weight.shape[0]:
                # gather + add + LayerNorm(128) in one kernel; its backward also does the three embedding backwards
                x = _emb.embed_layer_norm(self, input_ids, token_type_ids)
                return F.dropout(x, self.dropout_p, self.training)
        pos = torch.arange(S, device=input_ids.device)
        x = self.word_embeddings(input_ids) + self.token_type_embeddings(token_type_ids) \
            + self.position_embeddings(pos)[None]
        x = self.LayerNorm(x)
        return F.dropout(x, self.dropout_p, self.training)


class AlbertAttention(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        H = cfg.hidden_size
        self.num_heads, self.head_dim = cfg.num_attention_heads, H // cfg.num_attention_heads
        self.query, self.key, self.value, self.dense = _Linear(H, H), _Linear(H, H), _Linear(H, H), _Linear(H, H)
        self.LayerNorm = nn.LayerNorm(H, eps=cfg.layer_norm_eps)
        self.attn_dropout_p, self.out_dropout_p = cfg.attention_probs_dropout_prob, cfg.hidden_dropout_prob

    def forward(self, x, attn_bias, impl: str, fused_qkv=None, key_mask=None):
        B, S, H = x.shape
        if fused_qkv is not None:
            if (impl == "native" and self.head_dim == 64 and S % 128 == 0 and 128 <= S <= 512
                    and not ((self.attn_dropout_p > 0 or self.out_dropout_p > 0) and self.training)
                    and _native_ok(x, fused_qkv, impl) and _native_ok(x, self.dense, impl)):
                # the whole sub-layer as one autograd node: QKV GEMM -> tcgen05 attention on the packed projection ->
                # output GEMM (+ bias + residual); backward folds the residual-branch gradient into a GEMM epilogue
                from ..ops.attention import attention_block

                return _ln(attention_block(x, fused_qkv, self.dense, self.num_heads, key_mask), self.LayerNorm, impl)
            qkv = fused_linear(x, fused_qkv, impl)                        # [B,S,3H] in one GEMM
            if impl == "native" and self.head_dim == 64 and not (self.attn_dropout_p > 0 and self.training):
                from ..ops import attention as native_attention

                if native_attention.supported(qkv, self.num_heads):
                    # tcgen05 attention straight on the packed projection: no head transposes / unpack copies
                    ctx = native_attention.packed_attention(qkv, self.num_heads, key_mask)
                    if self.out_dropout_p > 0 and self.training:
                        y = F.dropout(fused_linear(ctx, self.dense, impl), self.out_dropout_p, True) + x
                    else:
                        y = fused_linear(ctx, self.dense, impl, residual=x)
                    return _ln(y, self.LayerNorm, impl)
            q, k, v = qkv.view(B, S, 3, self.num_heads, self.head_dim).permute(2, 0, 3, 1, 4)
        else:
            q = fused_linear(x, self.query, impl).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
            k = fused_linear(x, self.key, impl).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
            v = fused_linear(x, self.value, impl).view(B, S, self.num_heads, self.head_dim).transpose(1, 2)
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_bias,
                                             dropout_p=self.attn_dropout_p if self.training else 0.0)
        ctx = ctx.transpose(1, 2).reshape(B, S, H)
        if self.out_dropout_p > 0 and self.training:
            y = F.dropout(fused_linear(ctx, self.dense, impl), self.out_dropout_p, True) + x
        else:
            y = fused_linear(ctx, self.dense, impl, residual=x)          # residual fused in the epilogue
        return _ln(y, self.LayerNorm, impl)


class AlbertLayer(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.full_layer_layer_norm = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)
        self.attention = AlbertAttention(cfg)
        self.ffn = _Linear(cfg.hidden_size, cfg.intermediate_size)
        self.ffn_output = _Linear(cfg.intermediate_size, cfg.hidden_size)
        self.dropout_p = cfg.hidden_dropout_prob

    def forward(self, x, attn_bias, impl: str, fused_qkv=None, key_mask=None):
        a = self.attention(x, attn_bias, impl, fused_qkv, key_mask)
        if (_native_ok(a, self.ffn, impl) and self.ffn_output.out_features % 64 == 0
                and _main_grad(self.ffn_output.weight) is not None and not (self.dropout_p > 0 and self.training)):
            from ..ops.gemm import ffn as native_ffn

            return _ln(native_ffn(a, self.ffn, self.ffn_output, residual=a), self.full_layer_layer_norm, impl)
        h = fused_linear(a, self.ffn, impl, act="gelu")                  # gelu_new fused in the epilogue
        if self.dropout_p > 0 and self.training:
            y = F.dropout(fused_linear(h, self.ffn_output, impl), self.dropout_p, True) + a
        else:
            y = fused_linear(h, self.ffn_output, impl, residual=a)
        return _ln(y, self.full_layer_layer_norm, impl)


class AlbertLayerGroup(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.albert_layers = nn.ModuleList([AlbertLayer(cfg) for _ in range(cfg.inner_group_num)])


class AlbertEncoder(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.cfg = cfg
        self.embedding_hidden_mapping_in = _Linear(cfg.embedding_size, cfg.hidden_size)
        self.albert_layer_groups = nn.ModuleList([AlbertLayerGroup(cfg) for _ in range(cfg.num_hidden_groups)])


class _FusedQKV:
    """Per-forward fused view of the shared layer's query/key/value projections (one [3H,H] GEMM instead of
    three). The fused weight gradient lands in a private fp32 buffer that `flush()` scatters into the three
    parameters' main_grad buffers after backward."""

    def __init__(self, attn: AlbertAttention):
        self.attn = attn
        self.in_features = attn.query.in_features
        self.out_features = 3 * attn.query.out_features
        with torch.no_grad():
            self.weight = torch.cat([attn.query.weight, attn.key.weight, attn.value.weight], dim=0)
            self.weight.weight_t = self.weight.t().contiguous()     # K-major operand for the dgrad GEMM
            self.bias = torch.cat([attn.query.bias, attn.key.bias, attn.value.bias], dim=0)
        dev = self.weight.device
        self.weight.main_grad = torch.zeros(self.weight.shape, dtype=torch.float32, device=dev)
        self.bias.main_grad = torch.zeros(self.bias.shape, dtype=torch.float32, device=dev)

    @torch.no_grad()
    def flush(self) -> None:
        H = self.attn.query.out_features
        for i, lin in enumerate((self.attn.query, self.attn.key, self.attn.value)):
            lin.weight.main_grad.add_(self.weight.main_grad[i * H:(i + 1) * H])
            lin.bias.main_grad.add_(self.bias.main_grad[i * H:(i + 1) * H])


class AlbertModel(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.cfg = cfg
        self.embeddings = AlbertEmbeddings(cfg)
        self.encoder = AlbertEncoder(cfg)
        self.pooler = _Linear(cfg.hidden_size, cfg.hidden_size)

    def forward(self, input_ids, attention_mask, token_type_ids, impl: str, pending_flush: list):
        cfg = self.cfg
        x = self.embeddings(input_ids, token_type_ids, impl)
        x = fused_linear(x, self.encoder.embedding_hidden_mapping_in, impl)
        attn_bias = None
        if attention_mask is not None:
            # additive mask broadcast over heads and query positions, as HF's extended attention mask
            attn_bias = ((1.0 - attention_mask[:, None, None, :].to(x.dtype)) * torch.finfo(x.dtype).min)
        layers_per_group = cfg.num_hidden_layers // cfg.num_hidden_groups
        fused: Dict[int, _FusedQKV] = {}
        for i in range(cfg.num_hidden_layers):
            group = self.encoder.albert_layer_groups[i // layers_per_group]
            for j, layer in enumerate(group.albert_layers):
                fq = None
                use_fused = _native_ok(x, layer.attention.query, impl)
                if use_fused:
                    key = id(layer)
                    if key not in fused:
                        fused[key] = _FusedQKV(layer.attention)
                        with torch.no_grad():   # K-major W^T for the dgrad GEMMs, refreshed once per forward
                            for lin in (layer.attention.dense, layer.ffn, layer.ffn_output):
                                lin.weight.weight_t = lin.weight.detach().t().contiguous()
                        if torch.is_grad_enabled():
                            pending_flush.append(fused[key])
                    fq = fused[key]
                else:
                    for lin in (layer.attention.dense, layer.ffn, layer.ffn_output):
                        lin.weight.weight_t = None          # never leave a stale transposed copy behind
                x = layer(x, attn_bias, impl, fq, attention_mask)
        pooled = torch.tanh(fused_linear(x[:, 0], self.pooler, "torch"))
        return x, pooled


class AlbertMLMHead(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.LayerNorm = nn.LayerNorm(cfg.embedding_size, eps=cfg.layer_norm_eps)
        self.bias = nn.Parameter(torch.zeros(cfg.vocab_size))
        self.dense = _Linear(cfg.hidden_size, cfg.embedding_size)
        self.decoder = _Linear(cfg.embedding_size, cfg.vocab_size)
        self.decoder.bias = self.bias  # tied, as in HF

    def hidden(self, x, impl: str):
        h = fused_linear(x, self.dense, impl, act="gelu")
        if impl == "native" and h.is_cuda:
            from ..ops import embedding as _emb

            if _emb.ln128_supported(h, self.LayerNorm):
                return _emb.layer_norm128(h, self.LayerNorm)
        return self.LayerNorm(h)


class AlbertSOPHead(nn.Module):
    def __init__(self, cfg: AlbertConfig):
        super().__init__()
        self.dropout_p = cfg.classifier_dropout_prob
        self.classifier = _Linear(cfg.hidden_size, 2)

    def forward(self, pooled):
        return F.linear(F.dropout(pooled, self.dropout_p, self.training), self.classifier.weight,
                        self.classifier.bias)


@dataclass
class AlbertPreTrainingOutput:
    loss: Optional[torch.Tensor]
    mlm_loss: Optional[torch.Tensor] = None
    sop_loss: Optional[torch.Tensor] = None
    prediction_logits: Optional[torch.Tensor] = None
    sop_logits: Optional[torch.Tensor] = None


class AlbertForPreTraining(nn.Module):
    """forward(input_ids, attention_mask, token_type_ids, labels, sentence_order_label) -> .loss, like HF."""

    def __init__(self, cfg: AlbertConfig, impl: str = "torch", max_predictions_ratio: float = 0.25):
        super().__init__()
        self.config = cfg
        self.impl = impl
        self.max_predictions_ratio = max_predictions_ratio
        # masked positions that did not fit the static MLM capacity since the last read (0 with the 15 % recipe)
        self.register_buffer("mlm_dropped", torch.zeros((), dtype=torch.int64), persistent=False)
        self.albert = AlbertModel(cfg)
        self.predictions = AlbertMLMHead(cfg)
        self.sop_classifier = AlbertSOPHead(cfg)
        self._pending_flush: list = []
        self.apply(self._init_weights)
        self.tie_weights()

    def _init_weights(self, module):
        std = self.config.initializer_range
        if isinstance(module, _Linear):
            nn.init.normal_(module.weight, 0.0, std)
            nn.init.zeros_(module.bias)
        elif isinstance(module, nn.Embedding):
            nn.init.normal_(module.weight, 0.0, std)
            if module.padding_idx is not None:
                with torch.no_grad():
                    module.weight[module.padding_idx].zero_()
        elif isinstance(module, nn.LayerNorm):
            nn.init.ones_(module.weight)
            nn.init.zeros_(module.bias)

    def tie_weights(self):
        self.predictions.decoder.weight = self.albert.embeddings.word_embeddings.weight
        self.predictions.decoder.bias = self.predictions.bias

    def resize_token_embeddings(self, new_size: int):
        """Grow / shrink the vocabulary (albert/run_trainer.py:68)."""
        old = self.albert.embeddings.word_embeddings
        if new_size == old.num_embeddings:
            return old
        new = nn.Embedding(new_size, old.embedding_dim, padding_idx=old.padding_idx).to(old.weight)
        nn.init.normal_(new.weight, 0.0, self.config.initializer_range)
        n = min(new_size, old.num_embeddings)
        with torch.no_grad():
            new.weight[:n] = old.weight[:n]
            bias = torch.zeros(new_size, dtype=self.predictions.bias.dtype, device=self.predictions.bias.device)
            bias[:n] = self.predictions.bias[:n]
        self.albert.embeddings.word_embeddings = new
        self.predictions.bias = nn.Parameter(bias)
        self.predictions.decoder.out_features = new_size
        self.config.vocab_size = new_size
        self.tie_weights()
        return new

    # ------------------------------------------------------------------ forward
    def forward(self, input_ids, attention_mask=None, token_type_ids=None, labels=None,
                sentence_order_label=None, return_logits: bool = False) -> AlbertPreTrainingOutput:
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        seq, pooled = self.albert(input_ids, attention_mask, token_type_ids, self.impl, self._pending_flush)
        sop_logits = self.sop_classifier(pooled)
        B, S, H = seq.shape
        out = AlbertPreTrainingOutput(loss=None, sop_logits=sop_logits)
        dec_w = self.predictions.decoder.weight
        if return_logits or labels is None:
            h = self.predictions.hidden(seq, self.impl)
            out.prediction_logits = F.linear(h, dec_w, self.predictions.bias)
        if labels is not None:
            # gather the masked positions first (static capacity: no host sync, CUDA-graph friendly)
            flat_labels = labels.reshape(-1)
            is_masked = flat_labels != -100
            cap = min(B * S, max(64, int(math.ceil(B * S * self.max_predictions_ratio / 64.0)) * 64))
            order = torch.argsort(is_masked.to(torch.int8), descending=True, stable=True)[:cap]
            sel_labels = flat_labels[order]
            sel_valid = is_masked[order]
            h = self.predictions.hidden(seq.reshape(B * S, H)[order], self.impl)
            dec_b = self.predictions.bias
            if (self.impl == "native" and h.is_cuda and h.dtype == torch.bfloat16 and h.shape[0] % 64 == 0
                    and dec_w.shape[1] % 64 == 0 and dec_w.shape[0] % 8 == 0 and _main_grad(dec_w) is not None
                    and _main_grad(dec_b) is not None):
                # vocabulary projection on the tcgen05 GEMM as well (tied to the word embeddings: the weight gradient
                # accumulates into the embedding's fp32 main_grad; the dgrad reduces over V = 30000, whose tail of the
                # 64-wide K block is zero-filled by TMA)
                from ..ops.gemm import linear as native_linear

                logits = native_linear(h, dec_w, dec_b, wgrad_out=dec_w.main_grad, bgrad_out=dec_b.main_grad)
            else:
                logits = F.linear(h, dec_w, dec_b)
            if self.impl == "native" and _ce.supported(logits):
                # one pass for loss + log-sum-exp, one for the bf16 gradient: no fp32 copy of [cap, vocab]
                per_tok = _ce.cross_entropy_rows(logits, sel_labels.clamp_min(0))
            else:
                per_tok = F.cross_entropy(logits.float(), sel_labels.clamp_min(0), reduction="none")
            n_masked = is_masked.sum()
            # The capacity is static (no host sync, CUDA-graph friendly). Positions beyond it cannot be scored, so they
            # are taken out of the denominator too (the loss stays the mean over the scored positions) and counted on
            # the device: `mlm_dropped_positions()` reports them and the trainers warn when it is not zero.
            n_scored = sel_valid.sum()
            self.mlm_dropped += (n_masked - n_scored).to(self.mlm_dropped.dtype)
            denom = n_scored.clamp_min(1)
            mlm_loss = (per_tok * sel_valid).sum() / denom
            out.mlm_loss = mlm_loss
            out.loss = mlm_loss
            if sentence_order_label is not None:
                out.sop_loss = F.cross_entropy(sop_logits.float(), sentence_order_label.reshape(-1))
                out.loss = mlm_loss + out.sop_loss
        return out

    def mlm_dropped_positions(self, reset: bool = True) -> int:
        """Masked positions beyond `max_predictions_ratio` of the tokens that were left out of the loss (host read)."""
        n = int(self.mlm_dropped.item())
        if reset and n:
            self.mlm_dropped.zero_()
        return n

    def flush_fused_grads(self) -> None:
        """Scatter the fused-QKV weight gradients into the per-parameter main_grad buffers (after backward)."""
        for f in self._pending_flush:
            f.flush()
        self._pending_flush.clear()

    # ------------------------------------------------------------------ HF-compatible checkpoints
    def hf_state_dict(self) -> Dict[str, torch.Tensor]:
        sd = {k: v.detach().float().cpu() for k, v in self.state_dict().items()}
        sd["predictions.decoder.weight"] = sd["albert.embeddings.word_embeddings.weight"]
        sd["predictions.decoder.bias"] = sd["predictions.bias"]
        return sd

    def save_pretrained(self, directory: str) -> None:
        os.makedirs(directory, exist_ok=True)
        self.config.save_pretrained(directory)
        sd = self.hf_state_dict()
        try:
            from safetensors.torch import save_file

            # safetensors refuses aliased storage: tied tensors are stored once, HF re-ties them on load
            sd = {k: v.contiguous() for k, v in sd.items() if not k.startswith("predictions.decoder.")}
            save_file(sd, os.path.join(directory, "model.safetensors"), metadata={"format": "pt"})
        except ImportError:  # pragma: no cover
            torch.save(sd, os.path.join(directory, "pytorch_model.bin"))

    @classmethod
    def from_pretrained(cls, directory: str, impl: str = "torch", **kw) -> "AlbertForPreTraining":
        cfg = AlbertConfig.from_pretrained(directory)
        model = cls(cfg, impl=impl, **kw)
        st_path = os.path.join(directory, "model.safetensors")
        if os.path.exists(st_path):
            from safetensors.torch import load_file

            sd = load_file(st_path)
        else:
            sd = torch.load(os.path.join(directory, "pytorch_model.bin"), map_location="cpu")
        model.load_hf_state_dict(sd)
        return model

    def load_hf_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        # transformers >= 5 writes LayerNorm weights under their legacy TF names
        sd = {k.replace("LayerNorm.gamma", "LayerNorm.weight").replace("LayerNorm.beta", "LayerNorm.bias"): v
              for k, v in sd.items()}
        own = self.state_dict()
        missing = []
        with torch.no_grad():
            for k, v in own.items():
                if k in sd:
                    v.copy_(sd[k].to(v.dtype))
                elif k.startswith("predictions.decoder."):
                    continue
                else:
                    missing.append(k)
        if missing:
            raise KeyError(f"checkpoint is missing parameters: {missing[:5]}{'...' if len(missing) > 5 else ''}")
        self.tie_weights()

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())


class AlbertForTokenClassification(nn.Module):
    """Encoder + per-token linear classifier (HF names: `albert.*`, `classifier.*`); used by train_ner.py."""

    def __init__(self, cfg: AlbertConfig, num_labels: int, impl: str = "torch"):
        super().__init__()
        self.config, self.impl, self.num_labels = cfg, impl, num_labels
        self.albert = AlbertModel(cfg)
        self.dropout_p = cfg.classifier_dropout_prob
        self.classifier = _Linear(cfg.hidden_size, num_labels)
        self._pending_flush: list = []
        self.apply(lambda m: AlbertForPreTraining._init_weights(self, m))

    def forward(self, input_ids, attention_mask=None, token_type_ids=None, labels=None):
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        seq, _ = self.albert(input_ids, attention_mask, token_type_ids, self.impl, self._pending_flush)
        logits = F.linear(F.dropout(seq, self.dropout_p, self.training), self.classifier.weight, self.classifier.bias)
        loss = None
        if labels is not None:
            loss = F.cross_entropy(logits.float().view(-1, self.num_labels), labels.view(-1), ignore_index=-100)
        return AlbertPreTrainingOutput(loss=loss, prediction_logits=logits)

    def load_encoder_from(self, pretrained: "AlbertForPreTraining") -> None:
        self.albert.load_state_dict(pretrained.albert.state_dict())


class AlbertForSequenceClassification(nn.Module):
    """Encoder + pooled-output classifier (HF names: `albert.*`, `classifier.*`); used by train_ncc.py."""

    def __init__(self, cfg: AlbertConfig, num_labels: int, impl: str = "torch"):
        super().__init__()
        self.config, self.impl, self.num_labels = cfg, impl, num_labels
        self.albert = AlbertModel(cfg)
        self.dropout_p = cfg.classifier_dropout_prob
        self.classifier = _Linear(cfg.hidden_size, num_labels)
        self._pending_flush: list = []
        self.apply(lambda m: AlbertForPreTraining._init_weights(self, m))

    def forward(self, input_ids, attention_mask=None, token_type_ids=None, labels=None):
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        _, pooled = self.albert(input_ids, attention_mask, token_type_ids, self.impl, self._pending_flush)
        logits = F.linear(F.dropout(pooled, self.dropout_p, self.training), self.classifier.weight, self.classifier.bias)
        loss = F.cross_entropy(logits.float(), labels.view(-1)) if labels is not None else None
        return AlbertPreTrainingOutput(loss=loss, prediction_logits=logits)

    def load_encoder_from(self, pretrained: "AlbertForPreTraining") -> None:
        self.albert.load_state_dict(pretrained.albert.state_dict())
