"""Callable sub-modules with the reference's names, for callers that drive the model piecewise.

``src/translator.py:57-159`` does not call ``model.forward``: it calls ``model.ingredient_embeddings``, ``.forward_step``,
``.step_positional_encoding``, ``.step_wise_encoder``, ``.reasoner``, ``.Went``, ``.Wac``, ``.text_embeddings``,
``.decoder``, ``.pointer_generator_network`` and ``.decoder_classifier`` one video at a time, and ``src/train.py:691-707``
calls ``set_pretrained_embedding`` on three of them.  The nodes below give those names the reference's call
signatures and tensor shapes; each call runs the same CUDA operators as the batched step (``svpc_b200.model``) on a
one-video plan.  They exist for drop-in compatibility; the fast decoder is ``svpc_b200.decode.greedy_decode``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import functional as Fn
from .backend import ACT_RELU


class Node(nn.Module):
    """Plain container: holds parameters / buffers / child nodes under the reference's names."""

    def __init__(self, owner=None):
        super().__init__()
        object.__setattr__(self, "_owner", owner)

    @property
    def owner(self):
        return self._owner


def _ones_mask(n, device):
    return torch.ones(n, dtype=torch.float32, device=device)


class IngredientEmbeddings(Node):
    """BertEmbeddingsIngredientsUntied (reference model.py:516-537)."""

    def forward(self, input_ids, sep_masks):
        m = self.owner
        B = input_ids.shape[0]
        ni = [int(x) for x in sep_masks.reshape(B, -1).sum(dim=1).tolist()]
        plan = m.make_plan_raw([1] * B, ni, [0] * B, [{} for _ in range(B)], input_ids.device)
        E0 = m._ingredient_vectors(input_ids, sep_masks, plan)                      # (sum Ni, D) without padding rows
        D = E0.shape[1]
        out = torch.zeros(B, max(ni) if ni else 0, D, dtype=E0.dtype, device=E0.device)
        pe = m.p("ingredient_embeddings.position_embeddings_ingr.pe")
        out = out + pe[:out.shape[1]].to(out.dtype)                                # padded rows are 0 + pe (model.py:136-139)
        for b in range(B):
            out[b, :ni[b]] = E0[plan.cu_ingr[b]:plan.cu_ingr[b + 1]]
        return out

    def set_pretrained_embedding(self, pretrained_embedding, freeze=True):
        _set_word_embedding(self, pretrained_embedding, freeze)


class TextEmbeddings(Node):
    """BertEmbeddingsTextUntied (reference model.py:483-513)."""

    def forward(self, text_input_ids):
        m = self.owner
        S, Lt = text_input_ids.shape
        x = m._word_fc("text_embeddings", text_input_ids.reshape(-1).to(torch.int32),
                       pe=m.p("text_embeddings.position_embeddings_text.pe"), pe_period=Lt)
        return x.reshape(S, Lt, -1)

    def set_pretrained_embedding(self, pretrained_embedding, freeze=True):
        assert pretrained_embedding.shape == self.word_embeddings.weight.shape
        _set_word_embedding(self, pretrained_embedding, freeze)


def _set_word_embedding(node, pretrained, freeze):
    """nn.Embedding.from_pretrained(..., freeze, padding_idx) semantics on the node's ``word_embeddings.weight``."""
    w = node.word_embeddings.weight
    new = nn.Parameter(pretrained.detach().to(device=w.device, dtype=w.dtype).clone(), requires_grad=not freeze)
    node.word_embeddings._parameters["weight"] = new
    node.owner._invalidate()


class Reasoner(Node):
    """EntitiyReasoningNetwork (reference model.py:742-823); returns the reference's 5-tuple for ONE video."""

    def forward(self, step_vectors, ingr_vectors):
        m = self.owner
        V = step_vectors.reshape(-1, step_vectors.shape[-1])
        S, Ni = V.shape[0], ingr_vectors.shape[0]
        plan = m.make_plan_raw([S], [Ni], [0], [{}], V.device)
        prefix = "recipe_reasoner" if self is getattr(m, "recipe_reasoner", None) else "reasoner"
        ent, ac, ebar, Es, fbar = m._reasoner(prefix, V.to(m.compute_dtype), ingr_vectors.float().contiguous(), plan)
        return ent.view(S, Ni), ac, ebar, Es.view(S, Ni, -1), fbar

    def set_pretrained_embedding(self, pretrained_embedding, freeze):
        w = self.action_embeddings
        self._parameters["action_embeddings"] = nn.Parameter(
            pretrained_embedding.detach().to(device=w.device, dtype=w.dtype).clone(), requires_grad=not freeze)
        self.owner._invalidate()


class LinearReLU(Node):
    """nn.Sequential(nn.Linear, nn.ReLU): Went / Wac (reference model.py:861-864)."""

    def forward(self, x):
        m = self.owner
        shp = x.shape
        y = Fn.linear(x.reshape(-1, shp[-1]), getattr(self, "0").weight, getattr(self, "0").bias, act=ACT_RELU,
                      cd=m.compute_dtype)
        return y.reshape(*shp[:-1], -1)


class StepPositionalEncoding(Node):
    """PositionEncoding over a video's clips (reference model.py:80-104, 839)."""

    def forward(self, x):
        return x + self.pe[:x.shape[-2]].to(x.dtype)


class Encoder(Node):
    """BertEncoderNoMemoryUntied (reference model.py:594-617): ``encoder`` (clips) and ``step_wise_encoder``."""

    def forward(self, hidden_states, attention_mask, diagonal_mask=False, output_all_encoded_layers=True):
        m = self.owner
        assert not diagonal_mask, "the clip / step encoders are never causal (reference model.py:893, 1065)"
        N, L, D = hidden_states.shape
        prefix = "step_wise_encoder" if self is getattr(m, "step_wise_encoder", None) else "encoder"
        dev = hidden_states.device
        off = torch.arange(N, dtype=torch.int32, device=dev) * L
        ln = torch.full((N,), L, dtype=torch.int32, device=dev)
        lay = dict(q_off=off, q_len=ln, k_off=off, k_len=ln, max_lq=L, max_lk=L, D=D)
        x = hidden_states.reshape(N * L, D).to(m.compute_dtype)
        km = attention_mask.reshape(-1).to(torch.float32).contiguous()
        outs = []
        for l in range(m.config.num_hidden_layers):
            x = m._encoder_layer(x, f"{prefix}.layer.{l}", lay, km)
            outs.append(x.reshape(N, L, D))
        return outs if output_all_encoded_layers else outs[-1:]


class Decoder(Node):
    """BertDecoderNoMemoryUntied (reference model.py:666-694)."""

    def forward(self, hidden_states, attention_mask, ga_vectors, ga_masks, diagonal_mask=True,
                output_all_encoded_layers=False):
        m = self.owner
        assert diagonal_mask, "the caption decoder is always causal (reference translator.py:99, model.py:951)"
        S, Lt, D = hidden_states.shape
        C = ga_vectors.shape[1]
        plan = m.make_plan_raw([S], [0], [0], [{}], hidden_states.device)
        assert Lt == m.config.max_t_len
        x = hidden_states.reshape(S * Lt, D).to(m.compute_dtype)
        ctx = ga_vectors.reshape(S * C, D).to(m.compute_dtype)
        km = attention_mask.reshape(-1).to(torch.float32).contiguous()
        out = m._decoder(x, km, ctx, plan, C)
        return [out.reshape(S, Lt, D)]


class LMHead(Node):
    """BertLMPredictionHead (reference model.py:712-739) -> logits."""

    def forward(self, hidden_states):
        m = self.owner
        shp = hidden_states.shape
        return m._lm_logits(hidden_states.reshape(-1, shp[-1]).to(m.compute_dtype)).reshape(*shp[:-1], -1)


NODE_TYPES = {
    "ingredient_embeddings": IngredientEmbeddings,
    "text_embeddings": TextEmbeddings,
    "reasoner": Reasoner,
    "recipe_reasoner": Reasoner,
    "Went": LinearReLU,
    "Wac": LinearReLU,
    "step_positional_encoding": StepPositionalEncoding,
    "encoder": Encoder,
    "step_wise_encoder": Encoder,
    "decoder": Decoder,
    "decoder_classifier": LMHead,
}


def make_node(owner, path, name):
    return NODE_TYPES.get(path, Node)(owner)
