"""TEST INFRASTRUCTURE ONLY - CPU restatement of the four vocabulary-shard steps
(duodecoding_b200/csrc/shard_kernels.cuh) as a backend for ``ShardedVerifier``.

Used by tests/test_sharded_gloo.py to run the partition / all-gather / merge logic of
duodecoding_b200/sharded.py under gloo, world_size 2, without a GPU, and compare the result with the
unsharded oracle (oracle.fused_verify, which restates /root/reference/src/decoding_ops.py:109-220 and
src/utils.py:686-810).  Never imported by the product.
"""
from __future__ import annotations

import torch


class OracleShardBackend:
    def _job_rows(self, draft, target, gamma, t_temp, d_temp):
        """[B, 2*gamma+1, Vs] scaled logits in job order: draft j, target j, ..., bonus target."""
        rows = []
        for j in range(gamma):
            rows += [draft[:, j].float() / d_temp, target[:, j].float() / t_temp]
        rows.append(target[:, gamma].float() / t_temp)
        return torch.stack(rows, 1)

    def row_stats(self, draft, target, tokens, v_begin, t_temp, d_temp):
        B, gamma = tokens.shape
        x = self._job_rows(draft, target, gamma, t_temp, d_temp)
        Vs = x.shape[-1]
        mx = x.max(-1).values
        s = torch.exp(x - mx.unsqueeze(-1)).sum(-1)
        tokx, has = torch.zeros_like(mx), torch.zeros_like(mx)
        for j in range(gamma):
            loc = tokens[:, j] - v_begin
            inside = (loc >= 0) & (loc < Vs)
            idx = loc.clamp(0, Vs - 1).unsqueeze(-1)
            for k in (2 * j, 2 * j + 1):
                tokx[:, k] = torch.where(inside, x[:, k].gather(-1, idx).squeeze(-1), tokx[:, k])
                has[:, k] = inside.float()
        return torch.stack([mx, s, tokx, has], -1).reshape(B * (2 * gamma + 1), 4)

    @staticmethod
    def _merge(g):  # g [G, ..., 4] -> M, Z, tokx
        M = g[..., 0].max(0).values
        Z = (g[..., 1] * torch.exp(g[..., 0] - M)).sum(0)
        return M, Z, (g[..., 2] * g[..., 3]).sum(0)

    def accept(self, gathered, B, gamma, r):
        g = gathered.reshape(gathered.shape[0], B, 2 * gamma + 1, 4)
        M, Z, tokx = self._merge(g)
        prob = torch.exp(tokx - M) / Z
        sel_q, sel_p = prob[:, 0:2 * gamma:2], prob[:, 1:2 * gamma:2]
        mask = r <= sel_p / sel_q
        n = mask.to(torch.int8).cummin(1).values.sum(1).to(torch.int64) if gamma else torch.zeros(B, dtype=torch.int64)
        bonus = n == gamma
        prow = torch.where(bonus, torch.full_like(n, 2 * gamma), 2 * n + 1)
        qrow = (2 * n).clamp(max=2 * gamma)
        ar = torch.arange(B)
        plan = torch.stack([M[ar, prow], Z[ar, prow], M[ar, qrow], Z[ar, qrow], n.float(), bonus.float()], -1)
        out = {"accept_mask": mask, "selected_target_p": sel_p, "selected_draft_p": sel_q, "accepted_len": n,
               "all_accepted": bonus}
        return out, plan

    def sample_partial(self, draft, target, plan, noise, v_begin, t_temp, d_temp, gamma):
        B = target.shape[0]
        ar = torch.arange(B)
        n, bonus = plan[:, 4].long(), plan[:, 5] > 0
        p = torch.exp(target[ar, n].float() / t_temp - plan[:, 0:1]) / plan[:, 1:2]
        q = torch.exp(draft[ar, n.clamp(max=max(gamma - 1, 0))].float() / d_temp - plan[:, 2:3]) / plan[:, 3:4]
        d = torch.where(bonus.unsqueeze(-1), p, p - q)
        pos = d > 0
        mass = torch.where(pos, d, torch.zeros_like(d)).sum(-1)
        key = torch.where(pos, d / noise, torch.full_like(d, -1.0))
        kval, kidx = key.max(-1)          # first maximum = lowest index
        dz = torch.nan_to_num(d, nan=0.0, posinf=0.0, neginf=0.0)
        bval, bidx = dz.max(-1)
        return torch.stack([mass.double(), kval.double(), (kidx + v_begin).double(), bval.double(),
                            (bidx + v_begin).double()], -1)

    def finish(self, gathered, B):
        mass = gathered[..., 0].sum(0)
        def best(val, idx):  # max value, lowest global index on ties
            order = torch.stack([-val, idx], -1)            # [G, B, 2]
            pick = torch.zeros(B, dtype=torch.long)
            for g in range(1, order.shape[0]):
                cur = order[pick, torch.arange(B)]
                new = order[g]
                better = (new[:, 0] < cur[:, 0]) | ((new[:, 0] == cur[:, 0]) & (new[:, 1] < cur[:, 1]))
                pick = torch.where(better, torch.full_like(pick, g), pick)
            return val[pick, torch.arange(B)], idx[pick, torch.arange(B)]
        kval, kidx = best(gathered[..., 1], gathered[..., 2])
        _, bidx = best(gathered[..., 3], gathered[..., 4])
        use_key = (mass > 0) & (kval >= 0)
        return torch.where(use_key, kidx, bidx).long()
