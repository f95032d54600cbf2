"""TEST INFRASTRUCTURE: an fp64 numpy restatement of the llama / mixtral forward pass that build_llama describes
(/root/reference/src/llama.cpp:10453-10617; llm_build_kqv flash branch :9517-9575; llm_build_moe_ffn :9416-9515),
with weights dequantised by the oracle (bit-exact with the reference's dequantize_row_*).  Used to check whole-token
decode through the B200 backend: matmuls there run on q8_1-quantised activations, so agreement is within the
activation-quantisation noise (tolerances in the tests), not bit-exact."""
from __future__ import annotations

import numpy as np

import orc as O
from koboldcpp_b200.lib import F16, F32, Q8_0
from koboldcpp_b200.llama import ModelShape, tensor_shape


class RefModel:
    def __init__(self, shape: ModelShape, tensors: dict[str, tuple[int, np.ndarray]], n_ctx: int, type_kv: int = F16):
        self.s = shape
        self.w: dict[str, np.ndarray] = {}
        for name, (ty, raw) in tensors.items():
            ne0, ne1, ne2 = tensor_shape(shape, name)
            if ty == F32:
                a = raw.view(np.float32).astype(np.float64)
            else:
                a = O.dequantize(ty, raw.reshape(-1), ne0 * ne1 * ne2).astype(np.float64)
            self.w[name] = a.reshape(ne2, ne1, ne0) if ne2 > 1 else a.reshape(ne1, ne0)
        kvw = shape.head_dim * shape.n_head_kv
        self.k = np.zeros((shape.n_layer, n_ctx, kvw))
        self.v = np.zeros((shape.n_layer, n_ctx, kvw))
        self.type_kv = type_kv

    def _store(self, x: np.ndarray) -> np.ndarray:
        """What a cache row holds after the CPY node: f16 rounding, or the q8_0 / q4_0 round trip (quantize_row_q8_0_ref / q4_0_ref)."""
        if self.type_kv == F16:
            return x.astype(np.float32).astype(np.float16).astype(np.float64)
        rows = x.astype(np.float32)
        packed = O.quantize_rows(self.type_kv, rows)  # q8_0 or q4_0 blocks
        return O.dequantize(self.type_kv, packed.reshape(-1), rows.size).reshape(rows.shape).astype(np.float64)

    def _rms(self, x: np.ndarray, w: np.ndarray) -> np.ndarray:
        return x / np.sqrt((x * x).mean(axis=-1, keepdims=True) + self.s.rms_eps) * w

    def _rope(self, x: np.ndarray, pos: np.ndarray) -> np.ndarray:
        """LLAMA_ROPE_TYPE_NORM: pairs (2i, 2i+1), theta_i = pos * base^(-2i/D)  (ggml.c rope_yarn with ext_factor 0)."""
        T, H, D = x.shape
        i = np.arange(D // 2)
        theta = pos[:, None] * (self.s.rope_freq_base ** (-2.0 * i / D))[None, :]
        c, s_ = np.cos(theta)[:, None, :], np.sin(theta)[:, None, :]
        x0, x1 = x[..., 0::2], x[..., 1::2]
        out = np.empty_like(x)
        out[..., 0::2] = x0 * c - x1 * s_
        out[..., 1::2] = x0 * s_ + x1 * c
        return out

    def decode(self, tokens, n_past: int) -> np.ndarray:
        """Returns logits [n_tokens, n_vocab] (fp64) and updates the cache like llm_build_kv_store."""
        s = self.s
        T = len(tokens)
        D, H, HK = s.head_dim, s.n_head, s.n_head_kv
        pos = np.arange(n_past, n_past + T, dtype=np.float64)
        x = self.w["token_embd.weight"][np.asarray(tokens)]
        for il in range(s.n_layer):
            p = f"blk.{il}."
            h = self._rms(x, self.w[p + "attn_norm.weight"][0])
            q = (h @ self.w[p + "attn_q.weight"].T).reshape(T, H, D)
            k = (h @ self.w[p + "attn_k.weight"].T).reshape(T, HK, D)
            v = h @ self.w[p + "attn_v.weight"].T
            q = self._rope(q, pos)
            k = self._rope(k, pos).reshape(T, HK * D)
            self.k[il, n_past:n_past + T] = self._store(k)
            self.v[il, n_past:n_past + T] = self._store(v)
            n_kv = n_past + T
            K = self.k[il, :n_kv].reshape(n_kv, HK, D)
            V = self.v[il, :n_kv].reshape(n_kv, HK, D)
            rep = H // HK
            att = np.empty((T, H, D))
            for hh in range(H):
                kk, vv = K[:, hh // rep], V[:, hh // rep]
                sc = q[:, hh] @ kk.T / np.sqrt(D)
                for t in range(T):
                    sc[t, n_past + t + 1:] = -np.inf
                sc = sc - sc.max(axis=1, keepdims=True)
                pr = np.exp(sc)
                pr /= pr.sum(axis=1, keepdims=True)
                att[:, hh] = pr @ vv
            a = att.reshape(T, H * D) @ self.w[p + "attn_output.weight"].T
            ffn_inp = a + x
            h = self._rms(ffn_inp, self.w[p + "ffn_norm.weight"][0])
            if s.n_expert:
                logits = h @ self.w[p + "ffn_gate_inp.weight"].T
                pr = np.exp(logits - logits.max(axis=1, keepdims=True))
                pr /= pr.sum(axis=1, keepdims=True)
                out = np.zeros_like(h)
                for t in range(T):
                    sel = np.argsort(-pr[t], kind="stable")[: s.n_expert_used]
                    wts = pr[t, sel] / pr[t, sel].sum()
                    for e, wt in zip(sel, wts):
                        up = self.w[p + "ffn_up_exps.weight"][e] @ h[t]
                        g = self.w[p + "ffn_gate_exps.weight"][e] @ h[t]
                        act = g / (1.0 + np.exp(-g)) * up
                        out[t] += wt * (self.w[p + "ffn_down_exps.weight"][e] @ act)
                f = out
            else:
                up = h @ self.w[p + "ffn_up.weight"].T
                g = h @ self.w[p + "ffn_gate.weight"].T
                f = (g / (1.0 + np.exp(-g)) * up) @ self.w[p + "ffn_down.weight"].T
            x = f + ffn_inp
        h = self._rms(x, self.w["output_norm.weight"][0])
        return h @ self.w["output.weight"].T
