"""ORACLE (test infrastructure, not product code): CPU restatement of the CTC prefix scorer.

NumPy fp32 restatement of ``CTCPrefixScoreTH`` (``espnet/nets/ctc_prefix_score.py:10-220``)
for one utterance (``B = 1``, the only way espnet2 builds it, ``espnet/nets/scorers/ctc.py:96-98``)
and ``n`` running hypotheses.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module; the product path
(``llm_guided_asr_b200``) never does.

Parity pin: checked against outputs of the reference itself, recorded by
``tests/golden/make_golden.py`` (which imports ``/root/reference``) into ``tests/golden/*.npz``
(``tests/test_oracle_golden.py``).  The reference holds no golden vectors of its own for this
path.
"""

import numpy as np

LOGZERO = np.float32(-10000000000.0)  # ctc_prefix_score.py:32


def _lse2(a, b):
    """``torch.logsumexp`` of two fp32 values, ATen style: ``m + log(exp(a-m) + exp(b-m))``."""
    m = np.maximum(a, b)
    with np.errstate(under="ignore"):
        return (m + np.log(np.exp(a - m) + np.exp(b - m))).astype(np.float32)


def _lse_axis0(v):
    """``torch.logsumexp(v, dim=0)`` in fp32."""
    m = v.max(axis=0)
    with np.errstate(under="ignore"):
        return (m + np.log(np.exp(v - m[None]).sum(axis=0, dtype=np.float32))).astype(np.float32)


class CTCPrefixOracle:
    """One utterance: ``logp (T, V)`` fp32 log-posteriors, ``blank``, ``eos``."""

    def __init__(self, logp, blank, eos, xlen=None, margin=0):
        self.margin = margin  # attention-window margin (:56-61); 0 = no windowing
        x = np.array(logp, dtype=np.float32, copy=True)
        T, V = x.shape
        # padding rule of ctc_prefix_score.py:46-49
        if xlen is not None and xlen < T:
            x[xlen:, :] = LOGZERO
            x[xlen:, blank] = 0
        self.x = x
        self.T, self.V = T, V
        self.blank, self.eos = blank, eos
        self.end_frame = (T if xlen is None else xlen) - 1  # :54

    def initial_state(self):
        """``r_prev (T, 2, 1)``: r^n = logzero, r^b = cumsum_t x[t, blank]; ``s_prev = 0`` (:82-93)."""
        r = np.full((self.T, 2, 1), LOGZERO, dtype=np.float32)
        r[:, 1, 0] = np.cumsum(self.x[:, self.blank], dtype=np.float32)
        return r, np.zeros((1,), dtype=np.float32)

    def __call__(self, last_ids, output_length, r_prev, s_prev, ids, att_w=None, f_prev=(0, 1)):
        """One prefix-extension step (``CTCPrefixScoreTH.__call__`` :67-187).

        ``att_w (n, T)`` attention weights and ``f_prev = (f_min, f_max)`` of the previous call select the frame window
        when ``margin > 0`` (:143-148); ``self.last_window`` then holds ``(start, end, f_min, f_max)``.  The first call
        starts from ``(f_min, f_max) = (0, 1)`` (:93-94), so ``start >= 1`` always.

        last_ids (n,) int; output_length = len(y) - 1; r_prev (T, 2, n); s_prev (n,) prefix
        scores of the hypotheses; ids (n, S) candidate tokens or None (full vocabulary).
        Returns ``scores (n, V) = log_psi - s_prev`` and ``(r (T,2,n,S'), log_psi (n,V), idmap)``.
        """
        x, T, V = self.x, self.T, self.V
        n = len(last_ids)
        if ids is not None:
            S = ids.shape[1]
            idmap = np.full((n, V), -1, dtype=np.int64)  # :99-106
            for j in range(S):  # same write order as the advanced-index assignment
                idmap[np.arange(n), ids[:, j]] = j
            xn = x[:, ids.reshape(-1)].reshape(T, n, S)  # :108-113
        else:
            S = V
            idmap = None
            xn = np.broadcast_to(x[:, None, :], (T, n, V))
        xb = np.broadcast_to(x[:, self.blank][:, None, None], (T, n, S))

        r = np.full((T, 2, n, S), LOGZERO, dtype=np.float32)  # :122-127
        if output_length == 0:
            r[0, 0] = xn[0]  # :128-129

        r_sum = _lse2(r_prev[:, 0], r_prev[:, 1])  # (T, n)  :131
        log_phi = np.repeat(r_sum[:, :, None], S, axis=2)  # :132
        for h in range(n):  # :133-140
            if ids is not None:
                pos = idmap[h, last_ids[h]]
                if pos >= 0:
                    log_phi[:, h, pos] = r_prev[:, 1, h]
            else:
                log_phi[:, h, last_ids[h]] = r_prev[:, 1, h]

        if att_w is not None and self.margin > 0:  # :143-148
            f_arg = np.matmul(np.asarray(att_w, dtype=np.float32), np.arange(T, dtype=np.float32))
            f_min = max(int(f_arg.min()), f_prev[0])
            f_max = max(int(f_arg.max()), f_prev[1])
            start = min(f_prev[1], max(f_min - self.margin, output_length, 1))
            end = min(f_max + self.margin, T)
        else:
            f_min = f_max = 0
            start, end = max(output_length, 1), T  # :150-152
        self.last_window = (start, end, f_min, f_max)
        for t in range(start, end):  # :155-160
            rn = _lse2(r[t - 1, 0], log_phi[t - 1]) + xn[t]
            rb = _lse2(r[t - 1, 0], r[t - 1, 1]) + xb[t]
            r[t, 0], r[t, 1] = rn, rb

        # :163-178
        log_phi_x = np.concatenate((log_phi[0:1], log_phi[:-1]), axis=0) + xn
        cat = np.concatenate((log_phi_x[start:end], r[start - 1, 0][None]), axis=0)
        psi_c = _lse_axis0(cat)  # (n, S)
        if ids is not None:
            log_psi = np.full((n, V), LOGZERO, dtype=np.float32)
            for h in range(n):
                log_psi[h, ids[h]] = psi_c[h]
        else:
            log_psi = psi_c.copy()
        for h in range(n):  # :180-181 eos is always scored
            log_psi[h, self.eos] = r_sum[self.end_frame, h]
        if self.eos != self.blank:  # :183-185
            log_psi[:, self.blank] = LOGZERO
        scores = (log_psi - np.asarray(s_prev, dtype=np.float32).reshape(n, 1)).astype(np.float32)
        return scores, (r, log_psi, idmap)

    # -- streaming (ctc_prefix_score.py:222-269) ---------------------------------------------------
    def extend_prob(self, logp_new):
        """``extend_prob``: the utterance grew to ``logp_new (T_new, V)``; the frames already held keep
        their old posteriors (``self.x[:, :tmp_x.shape[1]] = tmp_x``, :239), the rest is appended."""
        xn = np.array(logp_new, dtype=np.float32, copy=True)
        if xn.shape[0] > self.T:
            xn[: self.T] = self.x
            self.x = xn
            self.T = xn.shape[0]
            self.end_frame = self.T - 1

    def extend_state(self, r_prev):
        """``extend_state``: ``r_prev (T_old, 2)`` -> ``(T, 2)``; new frames get r^n = logzero and
        r^b[t] = r^b[t-1] + x[t, blank], added one frame at a time (:263-267)."""
        r = np.full((self.T, 2), LOGZERO, dtype=np.float32)
        start = max(r_prev.shape[0], 1)
        r[:start] = r_prev[:start]
        for t in range(start, self.T):
            r[t, 1] = np.float32(r[t - 1, 1] + self.x[t, self.blank])
        return r

    @staticmethod
    def select_state(state, h, tok):
        """Survivor state (``CTCPrefixScorer.select_state`` TH branch, ``scorers/ctc.py:56-62``)."""
        r, log_psi, idmap = state
        col = tok if idmap is None else idmap[h, tok]  # -1 -> last column, as in the reference
        return r[:, :, h, col], log_psi[h, tok]
