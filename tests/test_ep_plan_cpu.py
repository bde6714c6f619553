"""The sender-side routing plan of the peer-memory EP path (``fused_ep.ep_plan``) and the padded grouped layout,
checked by simulating every rank on the CPU: pushes must land collision-free inside the right expert segment, the
receiver's tile table must cover exactly those rows, and a pull must bring every pair's row back."""
import pytest
import torch

from scaletorch_b200.ops.grouped import expert_counts
from scaletorch_b200.ops.grouped_native import GroupedLayout
from scaletorch_b200.parallel.expert_parallel.fused_ep import ep_plan


def test_expert_counts_matches_bincount():
    idx = torch.randint(0, 16, (100, 4))
    assert torch.equal(expert_counts(idx, 16), torch.bincount(idx.reshape(-1), minlength=16))
    assert expert_counts(torch.zeros(0, dtype=torch.long), 4).tolist() == [0, 0, 0, 0]


@pytest.mark.parametrize("ep,E,T,k,cf", [(2, 8, 50, 2, 2.0), (4, 16, 300, 4, 4.0), (4, 8, 64, 2, 0.5)])
def test_plan_is_collision_free_and_roundtrips(ep, E, T, k, cf):
    torch.manual_seed(ep * 100 + T)
    e_local = E // ep
    rows_cap = (int(cf * T * k) + e_local * 127 + 127) // 128 * 128
    idx = [torch.stack([torch.randperm(E)[:k] for _ in range(T)]) for _ in range(ep)]       # top-k: distinct experts
    x = [torch.randn(T, 8) for _ in range(ep)]
    table = torch.stack([expert_counts(i, E) for i in idx])                                    # what the exchange yields
    plans = [ep_plan(table, idx[r].reshape(-1), table[r], r, e_local, rows_cap) for r in range(ep)]
    # --- push: emulate the stores into every destination's receive region
    region = [torch.zeros(rows_cap, 8) for _ in range(ep)]
    owner = [torch.full((rows_cap,), -1, dtype=torch.long) for _ in range(ep)]                 # which expert wrote a row
    dropped = 0
    for r in range(ep):
        row, dest, n_drop, _, _ = plans[r]
        dropped += int(n_drop)
        flat = idx[r].reshape(-1)
        for p in range(T * k):
            if row[p] < 0:
                continue
            d, rw = int(dest[p]), int(row[p])
            assert owner[d][rw] == -1, "two pairs were sent to the same row"
            assert d == int(flat[p]) // e_local
            owner[d][rw] = int(flat[p])
            region[d][rw] = x[r][p // k]
    if cf >= ep:
        assert dropped == 0                                                                    # worst-case capacity
    # --- receiver layout: every written row lies in its expert's segment, padding rows stay zero
    for d in range(ep):
        _, _, _, offs_pad, tpe = plans[d]
        assert tpe.tolist() == table.sum(0).view(ep, e_local)[d].tolist()
        layout = GroupedLayout.from_padded_offsets(offs_pad, rows_cap, e_local)
        for e in range(e_local):
            lo, hi = int(offs_pad[e]), int(offs_pad[e + 1])
            assert lo % 128 == 0 or lo == rows_cap
            seg = owner[d][lo:hi]
            n = int((seg >= 0).sum())
            assert (seg[:n] == d * e_local + e).all() and (seg[n:] == -1).all()               # dense prefix, zero padding
            if cf >= ep:
                assert n == int(tpe[e])
        assert (owner[d][int(offs_pad[-1]):] == -1).all()
        te = layout.tile_expert.tolist()
        for t, e in enumerate(te):
            rows = owner[d][t * 128:(t + 1) * 128]
            if e < 0:
                assert (rows == -1).all()
            else:
                assert ((rows == -1) | (rows == d * e_local + e)).all()
    # --- pull with identity experts and unit weights: every token gets k copies of itself back (minus dropped pairs)
    for r in range(ep):
        row, dest, _, _, _ = plans[r]
        out = torch.zeros(T, 8)
        for p in range(T * k):
            if row[p] >= 0:
                out[p // k] += region[int(dest[p])][int(row[p])]
        kept = (row.view(T, k) >= 0).sum(1, keepdim=True).float()
        assert torch.allclose(out, x[r] * kept, atol=1e-5)


def test_grouped_layout_pad_unpad_roundtrip():
    counts = torch.tensor([5, 0, 130, 128, 1])
    R = int(counts.sum())
    layout = GroupedLayout(counts, R)
    x = torch.randn(R, 4)
    xp = layout.pad(x)
    assert xp.shape[0] == layout.r_max and layout.r_max % 128 == 0
    assert torch.equal(layout.unpad(xp), x)
    offs = layout.offs_pad.tolist()
    assert offs == [0, 128, 128, 384, 512, 640]
    te = layout.tile_expert.tolist()
    assert te[:5] == [0, 2, 2, 3, 4] and all(t == -1 for t in te[5:])
    assert float(xp.abs().sum()) == pytest.approx(float(x.abs().sum()), rel=1e-5)            # padding rows are zero
