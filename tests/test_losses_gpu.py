"""GPU parity of CriticalLoss (phoneme_to_articulation/principal_components/losses.py:23-99) and of
the closest-pair entry asb_min_distance, against the reference's golden numbers and the fp64 oracle."""
import os

import numpy as np
import pytest
import torch

from artspeech_b200 import synthetic

pytestmark = pytest.mark.gpu

ARTS = ["lower-lip", "pharynx", "soft-palate", "tongue", "upper-lip"]


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from artspeech_b200 import losses
    from oracle import geometry as ogeo
    return losses, ogeo


def test_min_distance_vs_oracle(mods):
    losses, ogeo = mods
    for n, m in ((50, 50), (15, 25), (20, 40), (7, 3), (1, 1), (33, 64)):
        u = torch.rand(257, 2, n)
        v = torch.rand(257, 2, m)
        v[0, :, m // 2] = u[0, :, n // 3]            # an exact zero
        d, ij = losses.min_distance(u.cuda(), v.cuda())
        full = ogeo._pairwise(u.double().transpose(-1, -2), v.double().transpose(-1, -2))   # (items, n, m)
        want = full.flatten(-2).min(dim=-1).values
        assert torch.allclose(d.cpu().double(), want, rtol=1e-5, atol=1e-12)
        # the reported pair attains the reported distance
        i, j = ij.cpu()[:, 0].long(), ij.cpu()[:, 1].long()
        got = full[torch.arange(257), i, j]
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-12)
        assert d[0].item() == 0.0
    # ties: the smallest flattened index i*m + j wins
    u = torch.zeros(1, 2, 4); v = torch.zeros(1, 2, 5)
    u[0, 0] = torch.tensor([0., 1., 0., 1.]); v[0, 0] = torch.tensor([5., 1., 0., 1., 0.])
    d, ij = losses.min_distance(u.cuda(), v.cuda())
    assert d.item() == 0.0 and ij.cpu().tolist() == [[0, 2]]
    d, ij = losses.min_distance(torch.zeros(0, 2, 4).cuda(), torch.zeros(0, 2, 5).cuda())
    assert d.shape == (0,) and ij.shape == (0, 2)


def test_critical_loss_golden_value_and_gradient(mods, golden_dir):
    losses, ogeo = mods
    g = np.load(os.path.join(golden_dir, "critical_loss.npz"))
    x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    mask = torch.from_numpy(g["mask"]).cuda()
    crit = losses.CriticalLoss(["LA", "TTCD", "TBCD", "VEL"], list(ARTS))
    loss = crit(x, x.detach(), torch.from_numpy(g["ref"]).cuda(), mask)
    loss.backward()
    rel = abs(loss.item() - float(g["loss64"])) / float(g["loss64"])
    print(f"critical loss {loss.item():.8f} vs reference fp64 {float(g['loss64']):.8f} (rel {rel:.1e}); "
          f"reference fp32 rel {abs(float(g['loss32']) - float(g['loss64'])) / float(g['loss64']):.1e}")
    assert rel <= 1e-5
    gerr = np.abs(x.grad.cpu().numpy() - g["grad64"]).max()
    assert gerr <= 1e-5 * np.abs(g["grad64"]).max(), gerr
    assert (x.grad.cpu().numpy() != 0).sum() == (g["grad64"] != 0).sum()
    # articulator list that already holds the upper incisor; a subset of the TVs
    x6 = torch.from_numpy(g["x6"]).cuda()
    loss6 = losses.CriticalLoss(["TTCD", "LA"], sorted(ARTS + ["upper-incisor"]))(x6, x6, None, mask[:, :2])
    assert abs(loss6.item() - float(g["loss6"])) / float(g["loss6"]) <= 1e-5
    # no TVs -> 0 (losses.py:59-61); denormalisation functions are applied per articulator (:80-82)
    assert losses.CriticalLoss([], list(ARTS))(x, x, None, mask).item() == 0
    den = {a: (lambda t: 2.0 * t) for a in ARTS}
    l2 = losses.CriticalLoss(["LA"], list(ARTS), denormalize_fn=den)(x.detach(), x.detach(),
                                                                     torch.from_numpy(g["ref"]).cuda(), mask[:, :1])
    l1 = losses.CriticalLoss(["LA"], list(ARTS))(x.detach(), x.detach(), torch.from_numpy(g["ref"]).cuda(),
                                                 mask[:, :1])
    assert abs(l2.item() - 2 * l1.item()) <= 1e-6 * l1.item()


def test_critical_loss_large_batch_vs_oracle(mods):
    losses, ogeo = mods
    B, T = 16, 100
    x = synthetic.contours(B * T, 5, 50, seed=31).view(B, T, 5, 2, 50)
    refarr = synthetic.UPPER_INCISOR_REF.view(1, 1, 1, 2, 50).expand(B, T, 1, 2, 50).contiguous()
    mask = (torch.rand(B, 4, T) < 0.3).long()
    want = ogeo.critical_loss(x, refarr, mask, ["LA", "TTCD", "TBCD", "VEL"], ARTS)
    got = losses.CriticalLoss(["LA", "TTCD", "TBCD", "VEL"], list(ARTS))(x.cuda(), x.cuda(), refarr.cuda(), mask.cuda())
    assert abs(got.item() - want.item()) / want.item() <= 1e-5
