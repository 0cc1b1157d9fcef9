"""BERT (base / large) for masked-LM + next-sentence pre-training.

BASELINE.json config 4: BERT-base pre-training with PairAveragingOptimizer.
The reference only knows BERT as a list of gradient sizes
(/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/model_sizes.py:27-81);
this is the architecture itself, bf16-friendly, attention through
torch.nn.functional.scaled_dot_product_attention (library flash kernels: model
math stays a library call, like in the reference)."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class BertConfig(object):
    def __init__(self, vocab_size=30522, hidden=768, layers=12, heads=12, intermediate=3072, max_pos=512,
                 type_vocab=2, dropout=0.1, eps=1e-12, fused_linear=False, fused_ln=False, fused_grads=False,
                 pad_vocab=1):
        self.vocab_size, self.hidden, self.layers, self.heads = vocab_size, hidden, layers, heads
        self.intermediate, self.max_pos, self.type_vocab, self.dropout, self.eps = intermediate, max_pos, type_vocab, dropout, eps
        # fused_linear: the FFN / projection layers run the tcgen05 kernel with the
        # bias (+GELU) epilogue (kungfu_b200.ops.linear); same parameters either way
        self.fused_linear = fused_linear
        # fused_ln: "add & norm" (dropout + residual + LayerNorm) as one kernel each way
        # (kungfu_b200.ops.fused_ln); same parameters either way
        self.fused_ln = fused_ln
        # fused_grads: bias gradients through the column-sum kernel and a one-pass gather for the packed
        # qkv gradient (kungfu_b200.ops.fused_ln) instead of autograd's generic reduce / stack / copy
        self.fused_grads = fused_grads
        # pad_vocab=8: embedding / decoder rows rounded up to a multiple of 8 so the 30522-way
        # decoder GEMMs take the aligned tensor-core kernels (pad logits are pinned to -1e4)
        self.pad_vocab = max(1, int(pad_vocab))
        self.vocab_rows = (vocab_size + self.pad_vocab - 1) // self.pad_vocab * self.pad_vocab


class _Layer(nn.Module):
    def __init__(self, c, index=0):
        super().__init__()
        from ..ops.linear import LinearAct
        self.heads = c.heads
        self.fused_ln = c.fused_ln
        self.fused_grads = c.fused_grads
        self.index = index
        # plain projections stay library GEMMs; the layer whose epilogue has work to fuse
        # (bias + GELU over the [tokens, intermediate] activation) runs the tcgen05 kernel
        self.qkv = nn.Linear(c.hidden, 3 * c.hidden)
        self.proj = nn.Linear(c.hidden, c.hidden)
        self.ln1 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.fc1 = LinearAct(c.hidden, c.intermediate, act='gelu', fused=c.fused_linear)
        self.fc2 = nn.Linear(c.intermediate, c.hidden)
        self.ln2 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.drop = c.dropout

    def _lin(self, m, x):
        """A plain projection: library GEMMs; with fused_grads its bias gradient comes from the
        single-launch column-sum kernel instead of a generic reduction."""
        if self.fused_grads:
            from ..ops.fused_ln import linear_fast_bias_grad
            return linear_fast_bias_grad(x, m.weight, m.bias)
        return m(x)

    def forward(self, x, mask=None):
        B, S, H = x.shape
        if self.fused_grads:
            from ..ops.fused_ln import split_qkv
            q, k, v = split_qkv(self._lin(self.qkv, x), self.heads)
        else:
            q, k, v = self.qkv(x).view(B, S, 3, self.heads, H // self.heads).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=self.drop if self.training else 0.0)
        a = a.transpose(1, 2).reshape(B, S, H)
        if self.fused_grads and not self.fc1.fused:
            ffn = lambda t: self._lin(self.fc2, F.gelu(self._lin(self.fc1, t)))  # noqa: E731
        else:
            ffn = lambda t: self._lin(self.fc2, self.fc1(t))  # noqa: E731  (fc1 applies the GELU)
        if self.fused_ln:
            from ..ops.fused_ln import add_dropout_layer_norm as adln
            x = adln(self._lin(self.proj, a), x, self.ln1.weight, self.ln1.bias, self.drop, self.training, self.ln1.eps,
                     stream_id=2 * self.index + 1)
            return adln(ffn(x), x, self.ln2.weight, self.ln2.bias, self.drop, self.training, self.ln2.eps,
                        stream_id=2 * self.index + 2)
        if self.fused_grads:
            x = self.ln1(x + F.dropout(self._lin(self.proj, a), self.drop, self.training))
            return self.ln2(x + F.dropout(ffn(x), self.drop, self.training))
        x = self.ln1(x + F.dropout(self.proj(a), self.drop, self.training))
        h = self.fc2(self.fc1(x))  # fc1 applies the GELU (fused epilogue or F.gelu)
        return self.ln2(x + F.dropout(h, self.drop, self.training))


class BertForPreTraining(nn.Module):
    def __init__(self, config=None):
        super().__init__()
        c = config or BertConfig()
        self.config = c
        self.tok = nn.Embedding(c.vocab_rows, c.hidden)
        self.pos = nn.Embedding(c.max_pos, c.hidden)
        self.typ = nn.Embedding(c.type_vocab, c.hidden)
        self.ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.layers = nn.ModuleList([_Layer(c, i) for i in range(c.layers)])
        self.pooler = nn.Linear(c.hidden, c.hidden)
        from ..ops.linear import LinearAct
        self.mlm_dense = LinearAct(c.hidden, c.hidden, act='gelu', fused=c.fused_linear)
        self.mlm_ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.mlm_bias = nn.Parameter(torch.zeros(c.vocab_rows))
        self.nsp = nn.Linear(c.hidden, 2)
        self.apply(self._init)
        if c.vocab_rows > c.vocab_size:
            with torch.no_grad():
                self.mlm_bias[c.vocab_size:] = -1e4  # pad tokens never win the softmax

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, std=0.02)
        if isinstance(m, nn.Linear) and m.bias is not None:
            nn.init.zeros_(m.bias)

    def forward(self, ids, token_type=None, mlm_labels=None, nsp_labels=None, mlm_positions=None):
        """``mlm_positions`` [B, P] (the original BERT input format: masked_lm_positions with
        mlm_labels [B, P] = masked_lm_ids) restricts the MLM head - a 30522-way projection - to the
        masked tokens; without it the head runs on every position and mlm_labels is [B, S] with -100
        at unmasked tokens. Same loss either way."""
        B, S = ids.shape
        pos = torch.arange(S, device=ids.device).unsqueeze(0)
        x = self.tok(ids) + self.pos(pos) + self.typ(token_type if token_type is not None else torch.zeros_like(ids))
        x = F.dropout(self.ln(x), self.config.dropout, self.training)
        if self.config.fused_ln and self.training and x.is_cuda:
            from ..ops.fused_ln import DropoutState
            DropoutState.advance(x.device)  # fresh dropout masks for this step (one tiny device op, graph-capturable)
        for layer in self.layers:
            x = layer(x)
        x_mlm = x
        if mlm_positions is not None:
            x_mlm = torch.gather(x, 1, mlm_positions.unsqueeze(-1).expand(-1, -1, x.shape[-1]))
        if self.config.fused_ln:
            from ..ops.fused_ln import add_dropout_layer_norm as adln
            h = adln(self.mlm_dense(x_mlm), None, self.mlm_ln.weight, self.mlm_ln.bias, 0.0, False, self.mlm_ln.eps)
        else:
            h = self.mlm_ln(self.mlm_dense(x_mlm))
        logits = F.linear(h, self.tok.weight, self.mlm_bias.to(h.dtype))  # tied decoder
        pooled = torch.tanh(self.pooler(x[:, 0]))
        nsp_logits = self.nsp(pooled)
        if mlm_labels is None:
            return logits[..., :self.config.vocab_size], nsp_logits
        loss = F.cross_entropy(logits.float().view(-1, logits.shape[-1]), mlm_labels.view(-1), ignore_index=-100)
        if nsp_labels is not None:
            loss = loss + F.cross_entropy(nsp_logits.float(), nsp_labels)
        return loss


def bert_base(**kw):
    """BERT-base for MLM + NSP pre-training (12 layers, hidden 768). Keyword arguments: any BertConfig field,
    e.g. fused_ln, fused_grads, fused_linear, pad_vocab, dropout."""
    return BertForPreTraining(BertConfig(**kw))


def bert_tiny(**kw):
    """2-layer toy configuration for tests."""
    kw = dict(dict(vocab_size=1000, hidden=128, layers=2, heads=2, intermediate=512, max_pos=128), **kw)
    return BertForPreTraining(BertConfig(**kw))


def bert_large(**kw):
    """BERT-large (24 layers, hidden 1024, 16 heads). Keyword arguments as for bert_base."""
    kw = dict(dict(hidden=1024, layers=24, heads=16, intermediate=4096), **kw)
    return BertForPreTraining(BertConfig(**kw))
