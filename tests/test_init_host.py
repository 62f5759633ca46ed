"""The initialiser kernel's random stream, replayed on the host (ops/init.py::host_random), and
the ParamSpec.dist tags that drive ``device_init``: every model family's tagged initialiser must
describe the same distribution as its host callable."""
import math

import pytest
import torch


def _philox(ctr, key):
    c, k = list(ctr), list(key)
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + 0x9E3779B9) & 0xFFFFFFFF, (k[1] + 0xBB67AE85) & 0xFFFFFFFF]
    return c


def test_philox_known_answers_and_host_stream(dtg):
    from dtg_b200.ops import init as di
    # Random123 known-answer vectors for Philox4x32-10
    assert _philox([0] * 4, [0] * 2) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert _philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert _philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]
    seed, sid = 0x1234567811223344, 77
    u = di.host_random(64, "uniform", 0.0, 1.0, seed=seed, stream_id=sid, first=5)
    want = [(_philox([e, 0, 0, sid], [seed & 0xFFFFFFFF, seed >> 32])[0] >> 8) / 16777216.0
            for e in range(5, 69)]
    assert u.tolist() == pytest.approx(want, abs=0)
    # element e is the same however the range is cut
    z = di.host_random(1000, "normal", 0.5, 2.0, seed=9)
    z2 = torch.cat([di.host_random(400, "normal", 0.5, 2.0, seed=9),
                    di.host_random(600, "normal", 0.5, 2.0, seed=9, first=400)])
    assert torch.equal(z, z2)
    assert not torch.equal(z, di.host_random(1000, "normal", 0.5, 2.0, seed=9, stream_id=1))
    n = 1 << 20
    zn = di.host_random(n, "normal", 0.0, 1.0, seed=1)
    assert abs(float(zn.mean())) < 5e-3 and abs(float(zn.std()) - 1.0) < 5e-3
    assert abs(float((zn ** 4).mean()) - 3.0) < 0.05
    t = di.host_random(n, "trunc_normal", 0.0, 1.0, seed=3)          # tf.truncated_normal
    assert float(t.abs().max()) <= 2.0 and abs(float(t.std()) - 0.8796) < 4e-3
    x = torch.empty(10, 7)
    di.init_random_(x, "uniform", -1.0, 1.0, seed=2)                 # CPU tensors take the host stream
    assert float(x.abs().max()) <= 1.0 and torch.equal(
        x.reshape(-1), di.host_random(70, "uniform", -1.0, 1.0, seed=2))


@pytest.mark.parametrize("family", ["mlp", "transformer", "transformer_post_ln", "resnet", "toy"])
def test_dist_tags_describe_the_host_initialisers(dtg, family):
    """ParamSpec.dist (what the device kernel draws) vs ParamSpec.init (what the host draws)."""
    from dtg_b200.models.mlp import MLP
    from dtg_b200.models.resnet import tiny_resnet
    from dtg_b200.models.toy import ToyModel
    from dtg_b200.models.transformer import tiny_transformer
    from dtg_b200.ops import init as di
    from dtg_b200.parallel.flat import FlatLayout
    model = {"mlp": lambda: MLP([784, 256, 10]), "transformer": tiny_transformer,
             "transformer_post_ln": lambda: tiny_transformer(post_ln=True),
             "resnet": tiny_resnet, "toy": lambda: ToyModel("normal", dim=4096)}[family]()
    specs = model.param_specs()
    assert all(s.dist is not None for s in specs)
    host = model.init_params(seed=0)
    for shards in (1, 3):
        layout = FlatLayout([(s.name, s.shape, s.shape) for s in specs], shards,
                            "balanced" if shards > 1 else "round_robin")
        flat = torch.full((layout.total,), 7.0)
        assert di.device_init(model, layout, flat, seed=11, host_fallback=False) == len(specs)
        got = {k: v.clone() for k, v in layout.unpack(flat).items()}
        if shards == 1:
            first = got
        else:                                    # same seed, same values, wherever a tensor landed
            for k in got:
                assert torch.equal(got[k], first[k]), k
    used = torch.zeros(layout.total, dtype=torch.bool)
    for e in layout.entries.values():
        used[e.offset:e.offset + e.numel] = True
    assert bool((flat[~used] == 0).all())        # alignment padding was zeroed, not left at 7
    for s in specs:
        a, b = got[s.name].float(), host[s.name].float()
        assert a.shape == b.shape
        kind, x = s.dist
        if kind == "constant":
            assert torch.equal(a, b) and bool((a == x).all())
        elif a.numel() >= 2048:                  # enough samples to compare moments
            assert float(a.std()) == pytest.approx(float(b.std()), rel=0.08), s.name
            assert abs(float(a.mean())) < 4 * float(b.std()) / math.sqrt(a.numel()) + 1e-6, s.name
            if kind == "uniform":
                assert float(a.abs().max()) <= x
            if kind == "trunc_normal":
                assert float(a.abs().max()) <= 2 * x
