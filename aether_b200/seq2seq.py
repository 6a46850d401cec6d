"""Decoder side of the reference's seq2seq Aether (nn/seq2seq/aether.py) on libaether_b200.so.

    RecurrentDecoder   nn/seq2seq/aether.py:505-654: same constructor (a ``params`` dict), parameter names, creation
                       order (seeded inits reproduce the reference's) and ``forward(inputs, hidden, edges,
                       predicted_field) -> (outputs, hidden)``; the step runs in csrc/seqdec.cu
    Seq2SeqAether      the decoder-side surface of the enclosing model (:14-191): ``decoder``, ``field_net``,
                       ``coordinate_embedding`` under the reference's names, ``predict_field``,
                       ``single_step_forward`` and ``predict_future``.  The edge-type ENCODER (:249-...) is not built:
                       pass the reference's own ``Encoder`` module as ``encoder=`` or give ``predict_future`` the sampled
                       edge types directly (``edges=``)
    DecoderRollout     T prediction steps (field predictor + decoder step each) captured in one CUDA graph

Inference only (no autograd through the kernels, dropout inactive); CUDA tensors only, no CPU / ATen fallback.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np
import torch
from torch import nn

from . import _lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class FourierFeatureMapper(nn.Module):
    """nn/nn/fourier_feature_mapper.py:7-23 (only the buffer lives here; the mapping runs in the field kernel)."""

    def __init__(self, in_size, out_size, std=1.0):
        super().__init__()
        rng = np.random.default_rng(42)
        self.register_buffer("B", torch.from_numpy(rng.normal(0, std, size=(in_size, out_size))).float())


class _NoWorkspacePickle:
    """torch.save(model) must not carry the scratch workspace (it is re-created on the next call)."""

    def __getstate__(self):
        d = self.__dict__.copy()
        d["_ws"] = None
        return d


class RecurrentDecoder(_NoWorkspacePickle, nn.Module):
    def __init__(self, params):
        super().__init__()
        self.num_vars = num_vars = params["num_vars"]
        input_size = params["input_size"]
        n_hid = params["decoder_hidden"]
        edge_types = params["num_edge_types"]
        self.dropout_prob = params["decoder_dropout"]
        self.edge_types = edge_types
        self.msg_fc1 = nn.ModuleList([nn.Linear(2 * n_hid, n_hid) for _ in range(edge_types)])
        self.msg_fc2 = nn.ModuleList([nn.Linear(n_hid, n_hid) for _ in range(edge_types)])
        self.msg_out_shape = n_hid
        self.skip_first_edge_type = params["skip_first"]
        self.hidden_r = nn.Linear(n_hid, n_hid, bias=False)
        self.hidden_i = nn.Linear(n_hid, n_hid, bias=False)
        self.hidden_h = nn.Linear(n_hid, n_hid, bias=False)
        self.present_r = nn.Linear(n_hid, n_hid, bias=True)
        self.present_i = nn.Linear(n_hid, n_hid, bias=True)
        self.present_n = nn.Linear(n_hid, n_hid, bias=True)
        self.out_mlp = nn.Sequential(
            nn.Linear(n_hid, n_hid), nn.ReLU(), nn.Dropout(p=self.dropout_prob),
            nn.Linear(n_hid, n_hid), nn.ReLU(), nn.Dropout(p=self.dropout_prob),
            nn.Linear(n_hid, input_size))
        self.send_edges, self.recv_edges = torch.where(~torch.eye(num_vars, dtype=torch.bool))
        self.use_3d = params.get("use_3d", False)
        self.num_dims = 3 if self.use_3d else 2
        if input_size != 2 * self.num_dims:
            raise ValueError("RecurrentDecoder: input_size must be 2 * num_dims (positions and velocities)")
        if edge_types > _lib.SEQ_MAX_TYPES:
            raise ValueError("RecurrentDecoder: at most %d edge types" % _lib.SEQ_MAX_TYPES)
        if n_hid % 4 != 0:
            raise ValueError("RecurrentDecoder: decoder_hidden must be a multiple of 4")
        self.num_orientations = self.num_dims * (self.num_dims - 1) // 2
        self.num_relative_features = 4 * self.num_dims + self.num_orientations
        self.num_pos_features = self.num_dims + self.num_orientations
        fe = 2 * self.num_relative_features + input_size + self.num_dims
        self.present_msg_fc1 = nn.ModuleList([nn.Linear(fe, n_hid) for _ in range(edge_types)])
        self.present_msg_fc2 = nn.ModuleList([nn.Linear(n_hid, n_hid) for _ in range(edge_types)])
        fr = input_size + self.num_relative_features + self.num_dims
        self.input_r = nn.Linear(fr, n_hid, bias=True)
        self.input_i = nn.Linear(fr, n_hid, bias=True)
        self.input_n = nn.Linear(fr, n_hid, bias=True)
        self._ws = None

    def get_initial_hidden(self, inputs):
        return torch.zeros(inputs.size(0), inputs.size(2), self.msg_out_shape, device=inputs.device)

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_normal_(m.weight.data)
                m.bias.data.fill_(0.1)

    # ---- C-ABI plumbing ----------------------------------------------------------------------------------------------
    def dims(self, batch: int, field_hidden: int = 0, rnn_hidden: int = 0, prior_hidden: int = 0) -> "_lib.AetherSeqDims":
        return _lib.AetherSeqDims(self.num_dims, self.num_vars, self.msg_out_shape, self.edge_types,
                                  1 if self.skip_first_edge_type else 0, field_hidden, rnn_hidden, prior_hidden,
                                  {"bf16": 1, "fp32x3": 2}.get(getattr(self, "precision", "fp32"), 0), batch)

    def c_params(self, field_net: Optional[nn.Module] = None, rff: Optional[torch.Tensor] = None):
        """(struct, tensors kept alive) for the library; every parameter must be a contiguous fp32 CUDA tensor."""
        p = _lib.AetherSeqParams()
        keep = []

        def dp(t):
            t = t.detach()
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError("aether_b200.seq2seq: parameters must be fp32 CUDA tensors (there is no CPU path)")
            t = t.contiguous()
            keep.append(t)
            return t.data_ptr()

        for i in range(self.edge_types):
            p.msg_fc1_w[i], p.msg_fc1_b[i] = dp(self.msg_fc1[i].weight), dp(self.msg_fc1[i].bias)
            p.msg_fc2_w[i], p.msg_fc2_b[i] = dp(self.msg_fc2[i].weight), dp(self.msg_fc2[i].bias)
            p.present_msg_fc1_w[i], p.present_msg_fc1_b[i] = dp(self.present_msg_fc1[i].weight), dp(self.present_msg_fc1[i].bias)
            p.present_msg_fc2_w[i], p.present_msg_fc2_b[i] = dp(self.present_msg_fc2[i].weight), dp(self.present_msg_fc2[i].bias)
        p.hidden_r_w, p.hidden_i_w, p.hidden_h_w = dp(self.hidden_r.weight), dp(self.hidden_i.weight), dp(self.hidden_h.weight)
        for n in ("r", "i", "n"):
            setattr(p, "present_%s_w" % n, dp(getattr(self, "present_" + n).weight))
            setattr(p, "present_%s_b" % n, dp(getattr(self, "present_" + n).bias))
            setattr(p, "input_%s_w" % n, dp(getattr(self, "input_" + n).weight))
            setattr(p, "input_%s_b" % n, dp(getattr(self, "input_" + n).bias))
        for j, k in enumerate((0, 3, 6)):
            p.out_w[j], p.out_b[j] = dp(self.out_mlp[k].weight), dp(self.out_mlp[k].bias)
        if field_net is not None:
            p.rff_B = dp(rff)
            for j, k in enumerate((0, 2, 4)):
                p.field_w[j], p.field_b[j] = dp(field_net[k].weight), dp(field_net[k].bias)
        return p, keep

    def workspace(self, dims, device) -> torch.Tensor:
        nbytes = _lib.load().aether_seq_workspace_bytes(ctypes.byref(dims))
        if nbytes == 0:
            raise RuntimeError("aether_b200.seq2seq: unsupported decoder shape")
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != device:
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return self._ws

    def _check(self, inputs, hidden, edges, field):
        B, N = inputs.shape[0], inputs.shape[1]
        E = N * (N - 1)
        if not inputs.is_cuda:
            raise RuntimeError("aether_b200.seq2seq: CUDA tensors only (there is no CPU path)")
        if N != self.num_vars or inputs.shape[2] != 2 * self.num_dims:
            raise RuntimeError("RecurrentDecoder: inputs must be [batch, %d, %d]" % (self.num_vars, 2 * self.num_dims))
        if tuple(hidden.shape) != (B, N, self.msg_out_shape):
            raise RuntimeError("RecurrentDecoder: hidden must be [batch, num_vars, decoder_hidden]")
        if tuple(edges.shape) != (B, E, self.edge_types):
            raise RuntimeError("RecurrentDecoder: edges must be [batch, num_vars*(num_vars-1), num_edge_types]")
        if tuple(field.shape) != (B, N, self.num_dims):
            raise RuntimeError("RecurrentDecoder: predicted_field must be [batch, num_vars, num_dims]")

    @torch.no_grad()
    def forward(self, inputs, hidden, edges, predicted_field):
        """inputs[B,N,2D], hidden[B,N,H], edges[B,N(N-1),types], predicted_field[B,N,D] -> (outputs, hidden)."""
        if self.training and self.dropout_prob > 0:
            raise RuntimeError("aether_b200.seq2seq.RecurrentDecoder is inference-only: call .eval() (dropout is not built)")
        self._check(inputs, hidden, edges, predicted_field)
        f32 = lambda t: t.detach().to(torch.float32).contiguous()
        inputs, hidden, edges, field = f32(inputs), f32(hidden), f32(edges), f32(predicted_field)
        lib = _lib.load()
        with torch.cuda.device(inputs.device):
            d = self.dims(inputs.shape[0])
            ws = self.workspace(d, inputs.device)
            p, keep = self.c_params()
            out, hid = torch.empty_like(inputs), torch.empty_like(hidden)
            _lib.check(lib.aether_seq_decoder_step(ctypes.byref(d), ctypes.byref(p), _lib.ptr(inputs), _lib.ptr(hidden),
                                                   _lib.ptr(edges), _lib.ptr(field), _lib.ptr(out), _lib.ptr(hid),
                                                   _lib.ptr(ws), ws.numel(), _stream()), "aether_seq_decoder_step")
        return out, hid

    @torch.no_grad()
    def localizer(self, extended_inputs):
        """AugmentedLocalizer.forward (nn/utils/augmented_global_to_local.py:51-68) on [B,N,3D] = [pos|vel|field]:
        (rel_feat, Rinv, edge_attr, edge_pos)."""
        D, N, F = self.num_dims, self.num_vars, self.num_relative_features
        x = extended_inputs.detach().to(torch.float32)
        states, field = x[..., :2 * D].contiguous(), x[..., 2 * D:].contiguous()
        B = x.shape[0]
        dev = x.device
        rel = torch.empty(B, N, 3 * D + F, device=dev)
        Rinv = torch.empty(B, N, D, D, device=dev)
        ea = torch.empty(B, N * (N - 1), 2 * F + 3 * D, device=dev)
        with torch.cuda.device(dev):
            d = self.dims(B)
            _lib.check(_lib.load().aether_seq_localizer(ctypes.byref(d), _lib.ptr(states), _lib.ptr(field), _lib.ptr(rel),
                                                        _lib.ptr(Rinv), _lib.ptr(ea), _stream()), "aether_seq_localizer")
        pos_idx = [2, 3, 4] if D == 2 else [3, 4, 5, 6, 7, 8]     # pos_representation='polar' (:18-24, :578-580)
        return rel, Rinv, ea, ea[..., pos_idx]


class RefNRIMLP(nn.Module):
    """Parameter holder with the reference's names (nn/utils/model_utils.py:15-54)."""

    def __init__(self, n_in, n_hid, n_out, do_prob=0.0, no_bn=False):
        super().__init__()
        self.model = nn.Sequential(nn.Linear(n_in, n_hid), nn.ELU(inplace=True), nn.Dropout(do_prob),
                                   nn.Linear(n_hid, n_out), nn.ELU(inplace=True))
        self.bn = None if no_bn else nn.BatchNorm1d(n_out)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_normal_(m.weight.data)
                m.bias.data.fill_(0.1)

    def bn_affine(self):
        """BatchNorm1d in eval mode as (scale, shift)."""
        if self.bn is None:
            return None, None
        scale = self.bn.weight.detach() / torch.sqrt(self.bn.running_var + self.bn.eps)
        return scale.contiguous(), (self.bn.bias.detach() - self.bn.running_mean * scale).contiguous()


class AnisotropicEdgeFilter(nn.Module):
    """Parameter holder with the reference's names (nn/nn/anisotropic_filter.py:12-33)."""

    def __init__(self, in_size, pos_size, hidden_size, out_size):
        super().__init__()
        self.num_relative_features, self.out_size = in_size, out_size
        self.edge_filter = nn.Sequential(nn.Linear(pos_size, hidden_size), nn.ELU(),
                                         nn.Linear(hidden_size, in_size * out_size))
        nn.init.orthogonal_(self.edge_filter[0].weight, gain=nn.init.calculate_gain("relu"))
        nn.init.orthogonal_(self.edge_filter[2].weight)


class Encoder(_NoWorkspacePickle, nn.Module):
    """Prior branch of the reference's edge-type encoder (nn/seq2seq/aether.py:249-425) on csrc/seqdec.cu: same
    constructor (``params`` dict), submodule / parameter names and creation order (a reference ``encoder.*`` state_dict
    loads strictly), ``forward(inputs, predicted_field)`` and ``single_step_forward(inputs, prior_state,
    predicted_field)`` as ``predict_future`` calls them.  Eval mode only (BatchNorm uses its running statistics, dropout
    is inactive).  The posterior branch (reverse LSTM + ``encoder_fc_out``, used by the training loss only) holds its
    parameters but is not computed: ``forward`` returns ``None`` in its place."""

    def __init__(self, params):
        super().__init__()
        num_vars = params["num_vars"]
        self.num_vars = num_vars
        self.num_edges = params["num_edge_types"]
        dropout = params["encoder_dropout"]
        hidden_size = params["encoder_hidden"]
        rnn_hidden_size = params["encoder_rnn_hidden"]
        if params.get("encoder_rnn_type", "lstm") != "lstm":
            raise ValueError("aether_b200.seq2seq.Encoder: only encoder_rnn_type='lstm' is built")
        inp_size = params["input_size"]
        self.mlp3 = RefNRIMLP(hidden_size, hidden_size, hidden_size, dropout)
        self.mlp4 = RefNRIMLP(hidden_size * 3, hidden_size, hidden_size, dropout)
        if rnn_hidden_size is None:
            rnn_hidden_size = hidden_size
        self.forward_rnn = nn.LSTM(hidden_size, rnn_hidden_size, batch_first=True)
        self.reverse_rnn = nn.LSTM(hidden_size, rnn_hidden_size, batch_first=True)

        def head(n_in, num_layers, hid):
            if num_layers == 1:
                return nn.Linear(n_in, self.num_edges)
            layers = [nn.Linear(n_in, hid), nn.ELU(inplace=True)]
            for _ in range(num_layers - 2):
                layers += [nn.Linear(hid, hid), nn.ELU(inplace=True)]
            layers.append(nn.Linear(hid, self.num_edges))
            return nn.Sequential(*layers)

        self.encoder_fc_out = head(2 * rnn_hidden_size, params["encoder_mlp_num_layers"], params.get("encoder_mlp_hidden"))
        self.prior_layers = params["prior_num_layers"]
        if not 1 <= self.prior_layers <= 4:
            raise ValueError("aether_b200.seq2seq.Encoder: prior_num_layers must be 1..4")
        self.prior_hidden = params.get("prior_hidden_size") if self.prior_layers > 1 else 0
        self.prior_fc_out = head(rnn_hidden_size, self.prior_layers, self.prior_hidden)
        self.use_3d = params.get("use_3d", False)
        self.num_dims = D = 3 if self.use_3d else 2
        if inp_size != 2 * D:
            raise ValueError("Encoder: input_size must be 2 * num_dims")
        if hidden_size % 4 != 0:
            raise ValueError("Encoder: encoder_hidden must be a multiple of 4")
        self.num_orientations = D * (D - 1) // 2
        self.num_relative_features = 4 * D + self.num_orientations
        self.num_pos_features = D + self.num_orientations
        self.res1 = nn.Linear(inp_size + self.num_relative_features + D, hidden_size)
        self.edge_filter = AnisotropicEdgeFilter(2 * self.num_relative_features + inp_size + D, self.num_pos_features,
                                                 hidden_size, hidden_size)
        self.pos_representation = params.get("pos_representation", "cart")
        if self.pos_representation not in ("cart", "polar"):
            raise ValueError
        self.hidden_size, self.rnn_hidden = hidden_size, rnn_hidden_size
        self.init_weights()
        self._ws = None

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_normal_(m.weight.data)
                m.bias.data.fill_(0.1)

    def _c_params(self):
        keep = []

        def dp(t):
            if t is None:
                return None
            t = t.detach()
            if not t.is_cuda or t.dtype != torch.float32:
                raise RuntimeError("aether_b200.seq2seq: parameters must be fp32 CUDA tensors (there is no CPU path)")
            t = t.contiguous()
            keep.append(t)
            return t.data_ptr()

        p = _lib.AetherSeqEncParams()
        f = self.edge_filter.edge_filter
        p.filter_w0, p.filter_b0, p.filter_w2, p.filter_b2 = dp(f[0].weight), dp(f[0].bias), dp(f[2].weight), dp(f[2].bias)
        p.res1_w, p.res1_b = dp(self.res1.weight), dp(self.res1.bias)
        for name in ("mlp3", "mlp4"):
            m = getattr(self, name)
            w, b = getattr(p, name + "_w"), getattr(p, name + "_b")
            w[0], b[0], w[1], b[1] = dp(m.model[0].weight), dp(m.model[0].bias), dp(m.model[3].weight), dp(m.model[3].bias)
            sc, sh = m.bn_affine()
            setattr(p, name + "_bn_scale", dp(sc))
            setattr(p, name + "_bn_shift", dp(sh))
        r = self.forward_rnn
        p.lstm_w_ih, p.lstm_w_hh, p.lstm_b_ih, p.lstm_b_hh = dp(r.weight_ih_l0), dp(r.weight_hh_l0), dp(r.bias_ih_l0), dp(r.bias_hh_l0)
        lins = [self.prior_fc_out] if self.prior_layers == 1 else [m for m in self.prior_fc_out if isinstance(m, nn.Linear)]
        for i, lin in enumerate(lins):
            p.prior_w[i], p.prior_b[i] = dp(lin.weight), dp(lin.bias)
        p.n_prior_layers = len(lins)
        p.pos_offset = 0 if self.pos_representation == "cart" else self.num_dims
        return p, keep

    def _dims(self, batch):
        return _lib.AetherSeqDims(self.num_dims, self.num_vars, 4, self.num_edges, 0, self.hidden_size, self.rnn_hidden,
                                  int(self.prior_hidden or 0), {"bf16": 1, "fp32x3": 2}.get(getattr(self, "precision", "fp32"), 0), batch)

    def initial_state(self, batch, device):
        n = batch * self.num_vars * (self.num_vars - 1)
        return (torch.zeros(1, n, self.rnn_hidden, device=device), torch.zeros(1, n, self.rnn_hidden, device=device))

    @torch.no_grad()
    def single_step_forward(self, inputs, prior_state, predicted_field):
        """inputs[B,N,2D], prior_state (h, c) each [1, B*N(N-1), R], predicted_field[B,N,D] -> (logits[B,E,types], state)."""
        if self.training:
            raise RuntimeError("aether_b200.seq2seq.Encoder is inference-only: call .eval()")
        if not inputs.is_cuda:
            raise RuntimeError("aether_b200.seq2seq: CUDA tensors only (there is no CPU path)")
        B, N = inputs.shape[0], inputs.shape[1]
        E = N * (N - 1)
        if N != self.num_vars or inputs.shape[2] != 2 * self.num_dims or tuple(predicted_field.shape) != (B, N, self.num_dims):
            raise RuntimeError("Encoder: inputs must be [batch, num_vars, 2*num_dims], the field [batch, num_vars, num_dims]")
        h, c = prior_state
        if h.numel() != B * E * self.rnn_hidden or c.numel() != h.numel():
            raise RuntimeError("Encoder: prior_state must hold [batch * num_edges, encoder_rnn_hidden] per tensor")
        f32 = lambda t: t.detach().to(torch.float32).contiguous()
        x, fld, h, c = f32(inputs), f32(predicted_field), f32(h), f32(c)
        lib = _lib.load()
        with torch.cuda.device(x.device):
            d = self._dims(B)
            nbytes = lib.aether_seq_workspace_bytes(ctypes.byref(d))
            if self._ws is None or self._ws.numel() < nbytes or self._ws.device != x.device:
                self._ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
            p, keep = self._c_params()
            logits = torch.empty(B, E, self.num_edges, device=x.device)
            h2, c2 = torch.empty_like(h), torch.empty_like(c)
            _lib.check(lib.aether_seq_encoder_step(ctypes.byref(d), ctypes.byref(p), _lib.ptr(x), _lib.ptr(fld), _lib.ptr(h),
                                                   _lib.ptr(c), _lib.ptr(logits), _lib.ptr(h2), _lib.ptr(c2),
                                                   _lib.ptr(self._ws), self._ws.numel(), _stream()), "aether_seq_encoder_step")
        return logits, (h2, c2)

    @torch.no_grad()
    def forward(self, inputs, predicted_field):
        """inputs[B,T,N,2D], predicted_field[B,N,T,D] -> (prior logits [B,T,E,types], None, final LSTM state)."""
        B, T = inputs.shape[0], inputs.shape[1]
        state = self.initial_state(B, inputs.device)
        out = []
        for t in range(T):
            lg, state = self.single_step_forward(inputs[:, t], state, predicted_field[:, :, t])
            out.append(lg)
        return torch.stack(out, 1), None, state


def gumbel_softmax(logits, tau=1.0, hard=False, eps=1e-10, generator=None):
    """nn/utils/model_utils.py:57-121 (noise drawn on the logits' device)."""
    U = torch.rand(logits.shape, device=logits.device, generator=generator)
    if hard and logits.is_cuda and logits.dtype == torch.float32 and eps == 1e-10:
        return sample_hard(logits, U)
    y_soft = torch.softmax((logits - torch.log(eps - torch.log(U + eps))) / tau, dim=-1)
    if not hard:
        return y_soft
    k = y_soft.argmax(-1, keepdim=True)
    return torch.zeros_like(y_soft).scatter_(-1, k, 1.0)


def sample_hard(logits, U, out=None):
    """One-hot arg-max of logits + gumbel(U) in one launch (aether_seq_sample_hard)."""
    lg, U = logits.detach().contiguous(), U.contiguous()
    out = torch.empty_like(lg) if out is None else out
    K = lg.shape[-1]
    with torch.cuda.device(lg.device):
        _lib.check(_lib.load().aether_seq_sample_hard(_lib.ptr(lg), _lib.ptr(U), _lib.ptr(out), lg.numel() // K, K, _stream()),
                   "aether_seq_sample_hard")
    return out


class Seq2SeqAether(nn.Module):
    def __init__(self, params, encoder: Optional[nn.Module] = None, precision: str = "fp32"):
        """precision: "fp32" (CUDA cores, 1e-5 against the reference) or "bf16" (every Linear on the tensor cores with
        bf16 operands and fp32 accumulation, 2e-2)."""
        super().__init__()
        if precision not in ("fp32", "bf16", "fp32x3"):
            raise ValueError("precision must be 'fp32', 'fp32x3' or 'bf16'")
        self.num_vars = params["num_vars"]
        self.encoder = encoder
        self.decoder = RecurrentDecoder(params)
        self.decoder.precision = precision
        if isinstance(encoder, Encoder):
            encoder.precision = precision
        self.num_edge_types = params.get("num_edge_types")
        self.gumbel_temp = params.get("gumbel_temp")
        self.use_3d = params.get("use_3d", False)
        self.num_dims = 3 if self.use_3d else 2
        hidden_size = params["encoder_hidden"]
        if hidden_size % 2 != 0:
            raise ValueError("Hidden size must be even")
        if hidden_size % 4 != 0:
            raise ValueError("Seq2SeqAether: encoder_hidden must be a multiple of 4")
        self.field_net = nn.Sequential(nn.Linear(hidden_size, hidden_size), nn.SiLU(), nn.Linear(hidden_size, hidden_size),
                                       nn.SiLU(), nn.Linear(hidden_size, self.num_dims))
        self.coordinate_embedding = FourierFeatureMapper(self.num_dims, hidden_size // 2, std=params.get("rff_std", 1.0))
        self.field_hidden = hidden_size

    def _dims(self, batch):
        return self.decoder.dims(batch, self.field_hidden)

    @classmethod
    def with_encoder(cls, params, precision: str = "fp32"):
        """The whole inference model: decoder side + this package's Encoder (prior branch)."""
        return cls(params, encoder=Encoder(params), precision=precision)

    def _params(self):
        return self.decoder.c_params(self.field_net, self.coordinate_embedding.B)

    @torch.no_grad()
    def predict_field(self, x):
        """nn/seq2seq/aether.py:83-87: (field at the positions in x[..., :D], the positions).  x: [..., N, >=D]."""
        coords = x[..., :self.num_dims]
        if x.shape[-1] < 2 * self.num_dims:
            raise RuntimeError("predict_field: states must carry positions and velocities")
        if not x.is_cuda:
            raise RuntimeError("aether_b200.seq2seq: CUDA tensors only (there is no CPU path)")
        xs = x.detach().to(torch.float32)[..., :2 * self.num_dims].reshape(-1, 2 * self.num_dims).contiguous()
        dec = self.decoder
        rows, N = xs.shape[0], dec.num_vars
        if rows % N != 0:
            raise RuntimeError("predict_field: the leading axes must hold whole trajectories of num_vars particles")
        field = torch.empty(rows, self.num_dims, device=xs.device)
        with torch.cuda.device(xs.device):
            d = self._dims(rows // N)
            ws = dec.workspace(d, xs.device)
            p, keep = self._params()
            _lib.check(_lib.load().aether_seq_predict_field(ctypes.byref(d), ctypes.byref(p), _lib.ptr(xs), _lib.ptr(field),
                                                            _lib.ptr(ws), ws.numel(), _stream()), "aether_seq_predict_field")
        return field.reshape(*x.shape[:-1], self.num_dims), coords

    def calculate_loss(self, *args, **kwargs):
        """nn/seq2seq/aether.py:102-153 (training: teacher forcing, NLL + KL) is not built: the kernels are inference-only."""
        raise RuntimeError("aether_b200.seq2seq.Seq2SeqAether.calculate_loss: seq2seq training is not built (inference only: "
                           "predict_future / single_step_forward); train with the reference and load its state_dict here")

    def single_step_forward(self, inputs, decoder_hidden, edge_logits, hard_sample, predicted_field):
        """nn/seq2seq/aether.py:89-100."""
        old_shape = edge_logits.shape
        edges = gumbel_softmax(edge_logits.reshape(-1, self.num_edge_types), tau=self.gumbel_temp,
                               hard=hard_sample).view(old_shape)
        predictions, decoder_hidden = self.decoder(inputs, decoder_hidden, edges, predicted_field)
        return predictions, decoder_hidden, edges

    @torch.no_grad()
    def predict_future(self, inputs, prediction_steps, return_edges=False, return_everything=False, edges=None):
        """nn/seq2seq/aether.py:155-191.  With ``edges[B, T-1+steps, N(N-1), types]`` the sampled edge types of every
        step are taken from the caller; otherwise ``self.encoder`` (the reference's Encoder) produces the logits."""
        T = inputs.size(1)
        hidden = self.decoder.get_initial_hidden(inputs)
        all_predictions, all_edges = [], []
        if edges is None:
            if self.encoder is None:
                raise RuntimeError("predict_future: no encoder was given to Seq2SeqAether; pass edges=")
            x = inputs[:, :-1].transpose(2, 1).contiguous()
            predicted_field, _ = self.predict_field(x)
            prior_logits, _, prior_hidden = self.encoder(inputs[:, :-1], predicted_field)
        for step in range(T - 1):
            cur = inputs[:, step]
            if edges is None:
                pred, hidden, e = self.single_step_forward(cur, hidden, prior_logits[:, step], True,
                                                           predicted_field[:, :, step])
            else:
                e = edges[:, step]
                pred, hidden = self.decoder(cur, hidden, e, self.predict_field(cur)[0])
            if return_everything:
                all_edges.append(e)
                all_predictions.append(pred)
        pred = inputs[:, T - 1]
        for step in range(prediction_steps):
            field, _ = self.predict_field(pred)
            if edges is None:
                logits, prior_hidden = self.encoder.single_step_forward(pred, prior_hidden, field)
                pred, hidden, e = self.single_step_forward(pred, hidden, logits, True, field)
            else:
                e = edges[:, T - 1 + step]
                pred, hidden = self.decoder(pred, hidden, e, field)
            all_predictions.append(pred)
            all_edges.append(e)
        predictions = torch.stack(all_predictions, dim=1)
        if return_edges:
            return predictions, torch.stack(all_edges, dim=1)
        return predictions


class DecoderRollout:
    """``steps`` prediction steps of predict_future's second loop (:176-184) with given edge types, as one CUDA graph:
    per step the field predictor and the decoder step (``aether_seq_rollout_step``, 14 launches)."""

    def __init__(self, model: Seq2SeqAether, batch: int, steps: int, cuda_graph: bool = True, device=None):
        self.model, self.steps, self.B = model, int(steps), int(batch)
        dec = model.decoder
        self.dev = torch.device(device) if device is not None else next(model.parameters()).device
        if self.dev.type != "cuda":
            raise RuntimeError("aether_b200.seq2seq: CUDA only (there is no CPU path)")
        N, D, H, K = dec.num_vars, dec.num_dims, dec.msg_out_shape, dec.edge_types
        f32 = dict(dtype=torch.float32, device=self.dev)
        self.x0 = torch.empty(batch, N, 2 * D, **f32)
        self.h0 = torch.zeros(batch, N, H, **f32)
        self.hidden = torch.empty(batch, N, H, **f32)
        self.edges = torch.empty(self.steps, batch, N * (N - 1), K, **f32)
        self.traj = torch.empty(self.steps, batch, N, 2 * D, **f32)
        self.dims = model._dims(batch)
        self.lib = _lib.load()
        self.ws = torch.empty(self.lib.aether_seq_workspace_bytes(ctypes.byref(self.dims)), dtype=torch.uint8, device=self.dev)
        self._use_graph = cuda_graph
        self._graph = None
        self.launches_per_replay = 0

    def _steps_sampled(self):
        """predict_future's second loop as the reference runs it (:176-184): field -> encoder step -> hard gumbel sample
        -> decoder step, every call on the current stream (capturable, torch's graph-safe generator draws the noise)."""
        m, dec, enc = self.model, self.model.decoder, self.model.encoder
        p, self._keep = m._params()
        pe, keep_e = enc._c_params()
        self._keep += keep_e
        de = enc._dims(self.B)
        self.hidden.copy_(self.h0)
        self.enc_h.copy_(self.enc_h0)
        self.enc_c.copy_(self.enc_c0)
        x = self.x0
        lib = self.lib
        for t in range(self.steps):
            _lib.check(lib.aether_seq_predict_field(ctypes.byref(self.dims), ctypes.byref(p), _lib.ptr(x), _lib.ptr(self.field),
                                                    _lib.ptr(self.ws), self.ws.numel(), _stream()), "aether_seq_predict_field")
            _lib.check(lib.aether_seq_encoder_step(ctypes.byref(de), ctypes.byref(pe), _lib.ptr(x), _lib.ptr(self.field),
                                                   _lib.ptr(self.enc_h), _lib.ptr(self.enc_c), _lib.ptr(self.logits),
                                                   _lib.ptr(self.enc_h), _lib.ptr(self.enc_c), _lib.ptr(self.ws_enc),
                                                   self.ws_enc.numel(), _stream()), "aether_seq_encoder_step")
            sample_hard(self.logits, torch.rand(self.logits.shape, device=self.dev), out=self.edges[t])
            _lib.check(lib.aether_seq_decoder_step(ctypes.byref(self.dims), ctypes.byref(p), _lib.ptr(x), _lib.ptr(self.hidden),
                                                   _lib.ptr(self.edges[t]), _lib.ptr(self.field), _lib.ptr(self.traj[t]),
                                                   _lib.ptr(self.hidden), _lib.ptr(self.ws), self.ws.numel(), _stream()),
                       "aether_seq_decoder_step")
            x = self.traj[t]

    @torch.no_grad()
    def run_sampled(self, x0, hidden0, prior_state) -> torch.Tensor:
        """The prediction loop with the model's own Encoder: x0[B,N,2D], the decoder's hidden state and the encoder's LSTM
        state after the burn-in -> predictions[steps,B,N,2D]; the sampled edge types are left in ``self.edges``."""
        enc = self.model.encoder
        if not isinstance(enc, Encoder):
            raise RuntimeError("DecoderRollout.run_sampled needs a model built with aether_b200.seq2seq.Encoder")
        with torch.cuda.device(self.dev):
            if not hasattr(self, "enc_h"):
                n = self.B * self.edges.shape[2]
                f32 = dict(dtype=torch.float32, device=self.dev)
                self.enc_h0, self.enc_c0 = torch.zeros(n, enc.rnn_hidden, **f32), torch.zeros(n, enc.rnn_hidden, **f32)
                self.enc_h, self.enc_c = torch.empty_like(self.enc_h0), torch.empty_like(self.enc_c0)
                self.logits = torch.empty(self.B, self.edges.shape[2], enc.num_edges, **f32)
                self.field = torch.empty(self.B, self.x0.shape[1], self.model.num_dims, **f32)
                de = enc._dims(self.B)
                self.ws_enc = torch.empty(self.lib.aether_seq_workspace_bytes(ctypes.byref(de)), dtype=torch.uint8, device=self.dev)
                self._graph_s = None
            self.x0.copy_(x0, non_blocking=True)
            self.h0.copy_(hidden0, non_blocking=True)
            self.enc_h0.copy_(prior_state[0].reshape(self.enc_h0.shape), non_blocking=True)
            self.enc_c0.copy_(prior_state[1].reshape(self.enc_c0.shape), non_blocking=True)
            if not self._use_graph:
                self._steps_sampled()
                return self.traj
            if self._graph_s is None:
                self._steps_sampled()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                n0 = self.lib.aether_launch_count()
                with torch.cuda.graph(g):
                    self._steps_sampled()
                self.launches_per_replay_sampled = int(self.lib.aether_launch_count() - n0)
                self._graph_s = g
            self._graph_s.replay()
            return self.traj

    def _steps(self):
        p, self._keep = self.model._params()
        self.hidden.copy_(self.h0)
        x = self.x0
        for t in range(self.steps):
            _lib.check(self.lib.aether_seq_rollout_step(ctypes.byref(self.dims), ctypes.byref(p), _lib.ptr(x),
                                                        _lib.ptr(self.hidden), _lib.ptr(self.edges[t]), _lib.ptr(self.traj[t]),
                                                        _lib.ptr(self.hidden), None, _lib.ptr(self.ws), self.ws.numel(),
                                                        _stream()), "aether_seq_rollout_step")
            x = self.traj[t]

    @torch.no_grad()
    def run(self, x0, edges, hidden0=None) -> torch.Tensor:
        """x0[B,N,2D] (the last burn-in state), edges[steps,B,N(N-1),types] (or [B,E,types] for every step),
        hidden0[B,N,H] (zeros when None) -> predictions[steps,B,N,2D]; the final hidden state is in ``self.hidden``."""
        with torch.cuda.device(self.dev):
            self.x0.copy_(x0, non_blocking=True)
            self.edges.copy_(edges if edges.dim() == 4 else edges.unsqueeze(0).expand_as(self.edges), non_blocking=True)
            if hidden0 is None:
                self.h0.zero_()
            else:
                self.h0.copy_(hidden0, non_blocking=True)
            if not self._use_graph:
                self._steps()
                return self.traj
            if self._graph is None:
                self._steps()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                n0 = self.lib.aether_launch_count()
                with torch.cuda.graph(g):
                    self._steps()
                self.launches_per_replay = int(self.lib.aether_launch_count() - n0)
                self._graph = g
            self._graph.replay()
            return self.traj
