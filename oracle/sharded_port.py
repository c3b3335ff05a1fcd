"""TEST INFRASTRUCTURE ONLY - CPU restatement of the vocabulary-shard steps (incl. the top-p radix select)
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

    def accept(self, gathered, B, gamma, r, tokens=None, cuts=None):
        g = gathered.reshape(gathered.shape[0], B, 2 * gamma + 1, 4)
        M, Z, tokx = self._merge(g)
        has = g[..., 3].sum(0) > 0
        INF = float("inf")
        xs = torch.full_like(M, -INF)
        istar = torch.full(M.shape, 2 ** 31 - 1, dtype=torch.int64)
        if cuts is not None:  # nucleus rows: the kept mass normalises, a filtered-out token has probability 0
            c = cuts.reshape(B, 2 * gamma + 1, 4)
            xs, istar, Z = c[..., 0].float(), c[..., 1].long(), c[..., 2].float()
            tk = torch.zeros(B, 2 * gamma + 1, dtype=torch.int64)
            for j in range(gamma):
                tk[:, 2 * j] = tk[:, 2 * j + 1] = tokens[:, j]
            has = has & ((tokx > xs) | ((tokx == xs) & (tk <= istar)))
        prob = torch.where(has, torch.exp(tokx - M) / Z, torch.zeros_like(M))
        sel_q, sel_p = prob[:, 0:2 * gamma:2], prob[:, 1:2 * gamma:2]
        mask = r <= sel_p / sel_q
        n = mask.to(torch.int8).cummin(1).values.sum(1).to(torch.int64) if gamma else torch.zeros(B, dtype=torch.int64)
        bonus = n == gamma
        prow = torch.where(bonus, torch.full_like(n, 2 * gamma), 2 * n + 1)
        qrow = (2 * n).clamp(max=2 * gamma)
        ar = torch.arange(B)
        plan = torch.stack([M[ar, prow], Z[ar, prow], M[ar, qrow], Z[ar, qrow], n.float(), bonus.float(),
                            xs[ar, prow], istar[ar, prow].float(), xs[ar, qrow], istar[ar, qrow].float()], -1).double()
        plan[:, 7], plan[:, 9] = istar[ar, prow].double(), istar[ar, qrow].double()  # (exact: beyond float32's integers)
        out = {"accept_mask": mask, "selected_target_p": sel_p, "selected_draft_p": sel_q, "accepted_len": n,
               "all_accepted": bonus}
        return out, plan

    def sample_partial(self, draft, target, plan, noise, v_begin, t_temp, d_temp, gamma, filtered=False):
        B = target.shape[0]
        ar = torch.arange(B)
        n, bonus = plan[:, 4].long(), plan[:, 5] > 0
        pf = plan.float()
        xp = target[ar, n].float() / t_temp
        xq = draft[ar, n.clamp(max=max(gamma - 1, 0))].float() / d_temp
        p = torch.exp(xp - pf[:, 0:1]) / pf[:, 1:2]
        q = torch.exp(xq - pf[:, 2:3]) / pf[:, 3:4]
        if filtered:
            gi = (torch.arange(target.shape[-1]) + v_begin).unsqueeze(0)
            p = torch.where((xp > pf[:, 6:7]) | ((xp == pf[:, 6:7]) & (gi <= plan[:, 7:8].long())), p, torch.zeros_like(p))
            q = torch.where((xq > pf[:, 8:9]) | ((xq == pf[:, 8:9]) & (gi <= plan[:, 9:10].long())), q, torch.zeros_like(q))
        d = torch.where(bonus.unsqueeze(-1), p, p - q)
        pos = d > 0
        mass = torch.where(pos, d, torch.zeros_like(d)).sum(-1)
        key = torch.where(pos, d / noise, torch.full_like(d, -1.0))
        kval, kidx = key.max(-1)          # first maximum = lowest index
        dz = torch.nan_to_num(d, nan=0.0, posinf=0.0, neginf=0.0)
        bval, bidx = dz.max(-1)
        return torch.stack([mass.double(), kval.double(), (kidx + v_begin).double(), bval.double(),
                            (bidx + v_begin).double()], -1)

    # ---- top-p under sharding: the protocol of duodecoding_b200/csrc/shard_nucleus.cuh (radix select over the
    # order-preserving key, 8 bits per level, 40-bit fixed-point integer masses) restated with torch integers
    SCALE = float(2 ** 40)

    @staticmethod
    def _key(x):
        bits = (x.float() + 0.0).contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        return torch.where(bits >= 2 ** 31, (~bits) & 0xFFFFFFFF, bits | 2 ** 31)

    @staticmethod
    def _key_inv(key):
        bits = torch.where(key >= 2 ** 31, key & 0x7FFFFFFF, (~key) & 0xFFFFFFFF)
        bits = torch.where(bits >= 2 ** 31, bits - 2 ** 32, bits)
        return bits.to(torch.int32).view(torch.float32)

    def _mass(self, x, M):
        return torch.round(torch.exp(x.double() - M.double()) * self.SCALE).to(torch.int64)

    def nucleus_levels(self, target, t_temp, d_temp):
        return 4

    def nucleus_level(self, draft, target, gamma, v_begin, t_temp, d_temp, top_p, d_top_p, gstats, sel, level, hist=None):
        B = target.shape[0]
        per = 2 * gamma + 1
        x = self._job_rows(draft, target, gamma, t_temp, d_temp).reshape(B * per, -1)
        rows = x.shape[0]
        if level == 0:
            g = gstats.reshape(gstats.shape[0], rows, 4)
            M = g[..., 0].max(0).values
            is_t = torch.tensor([(k == 2 * gamma) or (k & 1) == 1 for k in range(per)] * B)
            tp = torch.where(is_t, torch.tensor(float(top_p)), torch.tensor(float(d_top_p)))
            sel = {"M": M, "top_p": tp, "state": torch.where(tp > 0, 0, 2), "prefix": torch.zeros(rows, dtype=torch.int64),
                   "above": torch.zeros(rows, dtype=torch.int64), "target": torch.zeros(rows, dtype=torch.int64)}
        key = self._key(x)
        shift = 24 - 8 * level
        match = (sel["state"] == 0).unsqueeze(-1) & ((key >> (shift + 8)) == sel["prefix"].unsqueeze(-1) if level else
                                                     torch.ones_like(key, dtype=torch.bool))
        bins = ((key >> shift) & 255) + 256 * torch.arange(rows).unsqueeze(-1)
        mass = torch.where(match, self._mass(x, sel["M"].unsqueeze(-1)), torch.zeros_like(key))
        hm = torch.zeros(rows * 256, dtype=torch.int64).scatter_add_(0, bins.reshape(-1), mass.reshape(-1))
        hc = torch.zeros(rows * 256, dtype=torch.int64).scatter_add_(0, bins.reshape(-1), match.reshape(-1).long())
        return sel, torch.stack([hm.reshape(rows, 256), hc.reshape(rows, 256)], -1)

    def nucleus_reduce(self, ghist, rank, B, gamma, top_p, d_top_p, sel, level, levels, quota=None):
        m, c = ghist[..., 0].sum(0).flip(-1), ghist[..., 1].sum(0).flip(-1)  # descending key order
        rows = m.shape[0]
        inc = m.cumsum(-1)
        total = inc[:, -1]
        if level == 0:
            tgt = torch.floor(sel["top_p"].double() * total.double()).to(torch.int64)
            sel["target"] = torch.where(sel["top_p"] >= 1.0, total, tgt)
        above, target = sel["above"].unsqueeze(-1), sel["target"].unsqueeze(-1)
        cross = (above + inc > target) & (above + inc - m <= target)
        active = sel["state"] == 0
        found = cross.any(-1)
        pos = cross.float().argmax(-1)
        ar = torch.arange(rows)
        keep_all = active & ~found
        sel["above"] = torch.where(keep_all, sel["above"] + total, torch.where(active, sel["above"] + (inc - m)[ar, pos], sel["above"]))
        sel["state"] = torch.where(keep_all, 1, sel["state"])
        go = active & found
        sel["prefix"] = torch.where(go, (sel["prefix"] << 8) | (255 - pos), sel["prefix"])
        if quota is None:
            quota = torch.full((rows,), -1, dtype=torch.int64)
        if level == levels - 1:
            xs = self._key_inv(sel["prefix"])
            f = self._mass(xs, sel["M"])
            n = c[ar, pos]
            t = torch.where(f > 0, (sel["target"] - sel["above"]) // f.clamp(min=1) + 1, n).clamp(max=n)
            cg = ghist[..., 1].flip(-1)[:, ar, pos]  # [G, rows] ties per rank
            before = cg[:rank].sum(0)
            q = (t - before).clamp(min=0).minimum(cg[rank])
            quota = torch.where(go & (t < n), q, torch.full_like(q, -1))
            sel["above"] = torch.where(go, sel["above"] + t * f, sel["above"])
        return sel, quota

    def nucleus_tie(self, draft, target, gamma, v_begin, t_temp, d_temp, sel, quota, levels, cand=None):
        B = target.shape[0]
        x = self._job_rows(draft, target, gamma, t_temp, d_temp).reshape(B * (2 * gamma + 1), -1)
        key = self._key(x)
        out = torch.full((x.shape[0],), 2 ** 31 - 1, dtype=torch.int64)
        for row in range(x.shape[0]):
            q = int(quota[row])
            if int(sel["state"][row]) != 0 or q < 0:
                continue
            ties = (key[row] == sel["prefix"][row]).nonzero().reshape(-1)
            out[row] = int(ties[q - 1]) + v_begin if q > 0 else -1
        return out

    def nucleus_close(self, sel, gstats, gcand, B, gamma, levels, cuts=None):
        rows = sel["M"].shape[0]
        g = gstats.reshape(gstats.shape[0], rows, 4)
        _, Z, _ = self._merge(g)
        st = sel["state"]
        Zk = torch.where(st == 2, Z.double(), sel["above"].double() / self.SCALE)
        xs = torch.where(st == 0, self._key_inv(sel["prefix"]).double(), torch.full((rows,), -float("inf"), dtype=torch.float64))
        istar = torch.where(st == 0, gcand.reshape(-1, rows).max(0).values, torch.full((rows,), 2 ** 31 - 1, dtype=torch.int64))
        return torch.stack([xs, istar.double(), Zk, st.double()], -1)

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
