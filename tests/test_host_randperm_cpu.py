"""eda_host_randperm_batch: the batch's frame shuffles from torch's global CPU generator in one native call must be
indistinguishable from the per-sequence ``torch.randperm`` calls of the reference (models.py:229-232): same
permutations, same generator state afterwards (so whatever draws next is unaffected)."""
import numpy as np
import pytest
import torch


@pytest.mark.parametrize("seed,lens,pre", [(1, [7, 500, 1, 0, 2000, 2, 333], 0), (777, [500] * 64, 0), (777, [500] * 64, 5),
                                           (5, [2000, 1999, 3, 624, 625, 1248], 3), (2**31 + 11, [1, 1, 2, 3], 1)])
def test_native_batch_equals_torch_randperm(seed, lens, pre):
    from speaker_diarization_b200 import ops
    torch.manual_seed(seed)
    if pre:
        torch.rand(pre)                              # somewhere inside the generator's stream
    ref = [torch.randperm(n) for n in lens]
    state_ref = torch.get_rng_state()
    tail_ref = torch.randn(5)
    torch.manual_seed(seed)
    if pre:
        torch.rand(pre)
    out = torch.full((len(lens), max(lens)), -1, dtype=torch.int32)
    rows = out.data_ptr() + 4 * out.stride(0) * np.arange(len(lens), dtype=np.uint64)
    assert ops.host_randperm_rows(lens, rows)
    for i, n in enumerate(lens):
        assert torch.equal(out[i, :n].long(), ref[i]) and bool((out[i, n:] == -1).all())
    assert torch.equal(torch.get_rng_state(), state_ref)
    assert torch.equal(torch.randn(5), tail_ref)


def test_model_shuffles_are_the_reference_draws():
    """_draw_perm / draw_perms go through the native call: compare with torch.randperm on a CPU-only model object."""
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    m = TransformerEdaModel(2, 345, 4, 256, 1)
    groups, order = [[2000, 2000, 2000], [800, 303]], [(0, 0), (0, 1), (1, 0), (0, 2), (1, 1)]
    torch.manual_seed(3)
    want = [torch.randperm(groups[g][i]) for g, i in order]
    torch.manual_seed(3)
    slots = [m._perm_slot(len(g), max(g)) for g in groups] if torch.cuda.is_available() else None
    if slots is None:                      # no GPU here: exercise the same arithmetic on plain host buffers
        from speaker_diarization_b200 import ops
        bufs = [torch.full((len(g), max(g)), -1, dtype=torch.int32) for g in groups]
        gi = np.asarray(order, dtype=np.int64)
        base = np.asarray([b.data_ptr() for b in bufs], dtype=np.uint64)
        stride = np.asarray([4 * max(g) for g in groups], dtype=np.uint64)
        rows = base[gi[:, 0]] + stride[gi[:, 0]] * gi[:, 1].astype(np.uint64)
        assert ops.host_randperm_rows([groups[g][i] for g, i in order], rows)
        for (g, i), w in zip(order, want):
            assert torch.equal(bufs[g][i, : groups[g][i]].long(), w)
