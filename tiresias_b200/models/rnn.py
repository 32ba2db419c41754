"""The trace's recurrent and BERT checkpoints as executable models (model/model_factory.py:19-29):
``2_layers_lstm_gigaword`` (330 MB), ``1_layer_bilstm_opennmt`` (900 MB), ``4_layers_brnn`` (1300 MB) — OpenNMT-style
sequence-to-sequence LSTMs with global dot attention — and ``BERT_Chinese`` (350 MB, a BERT-base encoder with the
21128-token Chinese vocabulary). Round 1 ran Transformers in their place; what the PS/worker job moves per
iteration is the gradient of THESE parameter sets: two large embedding tables, a generator matrix of vocabulary
size and comparatively small recurrent weights.

The recurrent cells run on cuDNN's fused LSTM (no hand-written recurrent kernel here); every projection
(attention output, generator, BERT's dense layers) runs on the tcgen05 GEMM through :class:`TcLinear`.
Next-token loss on synthetic tokens: the source is the batch, the target the batch shifted by one.
"""
from __future__ import annotations

import torch
from torch import nn
from torch.nn import functional as F

from ..ops import native
from ..ops.linear import TcLinear
from .transformer import Block, _lin, _ln


class Seq2SeqLSTM(nn.Module):
    def __init__(self, vocab: int, emb: int, hidden: int, layers: int, bidirectional: bool, fused: bool = True):
        super().__init__()
        self.vocab = vocab
        self.src_emb = nn.Embedding(vocab, emb)
        self.tgt_emb = nn.Embedding(vocab, emb)
        enc_hidden = hidden // 2 if bidirectional else hidden
        self.encoder = nn.LSTM(emb, enc_hidden, layers, batch_first=True, bidirectional=bidirectional)
        self.decoder = nn.LSTM(emb, hidden, layers, batch_first=True)
        self.attn_out = _lin(2 * hidden, hidden, fused, bias=False)          # Luong global attention, "dot" score
        self.generator = _lin(hidden, vocab, fused)
        self.layers, self.bidirectional = layers, bidirectional
        for e in (self.src_emb, self.tgt_emb):
            nn.init.normal_(e.weight, std=0.05)

    def _bridge(self, state: torch.Tensor) -> torch.Tensor:
        if not self.bidirectional:
            return state
        l2, b, h = state.shape                                  # [layers*2, B, H/2] -> [layers, B, H]
        return state.view(l2 // 2, 2, b, h).transpose(1, 2).reshape(l2 // 2, b, 2 * h).contiguous()

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        self.encoder.flatten_parameters()
        self.decoder.flatten_parameters()
        memory, (h, c) = self.encoder(self.src_emb(tokens))
        # teacher forcing: the decoder reads the batch itself and predicts its shift (the trainer's labels)
        dec, _ = self.decoder(self.tgt_emb(tokens), (self._bridge(h), self._bridge(c)))
        score = torch.bmm(dec, memory.transpose(1, 2))
        ctx = torch.bmm(F.softmax(score.float(), dim=-1).to(dec.dtype), memory)
        out = torch.tanh(self.attn_out(torch.cat((ctx, dec), dim=-1)))
        return self.generator(out)


def lstm_gigaword(fused: bool = True) -> Seq2SeqLSTM:
    """2-layer LSTM summariser: 2 x 50k x 500 embeddings + a 500 x 50k generator dominate (~81 M parameters,
    325 MB fp32 against the table's 330)."""
    return Seq2SeqLSTM(vocab=50000, emb=500, hidden=500, layers=2, bidirectional=False, fused=fused)


def bilstm_opennmt(fused: bool = True) -> Seq2SeqLSTM:
    """1-layer bidirectional encoder, 1-layer decoder, hidden 1024, 70k vocabulary (~224 M parameters, 895 MB)."""
    return Seq2SeqLSTM(vocab=70000, emb=1024, hidden=1024, layers=1, bidirectional=True, fused=fused)


def brnn_4_layers(fused: bool = True) -> Seq2SeqLSTM:
    """4-layer bidirectional encoder / 4-layer decoder, hidden 1024, 90k vocabulary (~327 M parameters, 1.3 GB)."""
    return Seq2SeqLSTM(vocab=90000, emb=1024, hidden=1024, layers=4, bidirectional=True, fused=fused)


class BertEncoder(nn.Module):
    """BERT-base (12 x 768, 12 heads, FFN 3072) with the Chinese vocabulary (21128) and an LM head tied to the
    token embedding: ~102 M parameters (407 MB fp32; the table's BERT_Chinese: 350). Bidirectional attention; the
    synthetic objective is the trainer's next-token loss over the same positions."""

    def __init__(self, vocab: int = 21128, d_model: int = 768, n_layers: int = 12, n_heads: int = 12,
                 d_ff: int = 3072, max_len: int = 512, fused: bool = True):
        super().__init__()
        self.vocab = vocab
        self.tok = nn.Embedding(vocab, d_model)
        self.pos = nn.Embedding(max_len, d_model)
        self.seg = nn.Embedding(2, d_model)
        self.ln_e = _ln(d_model, fused)
        self.blocks = nn.ModuleList(Block(d_model, n_heads, d_ff, fused, causal=False) for _ in range(n_layers))
        self.ln_f = _ln(d_model, fused)
        self.head = _lin(d_model, vocab, fused, bias=False)
        self.head.weight = self.tok.weight                   # tied, as in BERT's masked-LM head
        for e in (self.tok, self.pos, self.seg):
            nn.init.normal_(e.weight, std=0.02)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        b, s = tokens.shape
        x = self.tok(tokens) + self.pos(torch.arange(s, device=tokens.device)) + self.seg.weight[0]
        x = self.ln_e(x)
        for blk in self.blocks:
            x = blk(x)
        return self.head(self.ln_f(x))


def bert_chinese(fused: bool = True) -> BertEncoder:
    return BertEncoder(fused=fused)
