"""Zero-DCE checkpoint import: a torch module with the public model's layer names and
forward pass (written here from the paper's description: seven 3x3 convs, skip concats, tanh, eight applications of
x + r (x^2 - x)) is the independent reference; its state_dict goes through load_dce_state_dict() and the oracle /
the GPU path must reproduce its output with curve_sign = -1.  No reference file exists to cite
(/root/reference/README.md:1-2)."""
import numpy as np
import pytest
import torch

import oracle
import low_light_image_enhancement_b200 as L


class PublicDceNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        nf = 32
        conv = lambda i, o: torch.nn.Conv2d(i, o, 3, 1, 1, bias=True)
        self.e_conv1, self.e_conv2, self.e_conv3, self.e_conv4 = conv(3, nf), conv(nf, nf), conv(nf, nf), conv(nf, nf)
        self.e_conv5, self.e_conv6, self.e_conv7 = conv(2 * nf, nf), conv(2 * nf, nf), conv(2 * nf, 24)

    def forward(self, x):
        relu = torch.relu
        x1 = relu(self.e_conv1(x)); x2 = relu(self.e_conv2(x1)); x3 = relu(self.e_conv3(x2)); x4 = relu(self.e_conv4(x3))
        x5 = relu(self.e_conv5(torch.cat([x3, x4], 1)))
        x6 = relu(self.e_conv6(torch.cat([x2, x5], 1)))
        r = torch.tanh(self.e_conv7(torch.cat([x1, x6], 1)))
        for ri in torch.split(r, 3, dim=1):
            x = x + ri * (x * x - x)
        return x, r


def _net(seed=0):
    torch.manual_seed(seed)
    net = PublicDceNet().double()
    with torch.no_grad():
        for m in net.modules():
            if isinstance(m, torch.nn.Conv2d):
                m.weight.normal_(0.0, (2.0 / (9 * m.in_channels)) ** 0.5)
                m.bias.uniform_(-0.05, 0.05)
    return net


def test_state_dict_to_blob_matches_public_forward(tmp_path):
    net = _net()
    x = oracle.synth_lowlight(40, 56, 3)
    with torch.no_grad():
        want, r = net(torch.from_numpy(x.transpose(2, 0, 1)[None]).double())
    sd = {k: v.float() for k, v in net.state_dict().items()}
    path = tmp_path / "Epoch99.pth"
    torch.save({"module." + k: v for k, v in sd.items()}, path)          # DataParallel-style prefix, from disk
    blob = L.load_dce_state_dict(path)
    assert blob.shape == (79416,) and np.array_equal(blob, L.load_dce_state_dict({"state_dict": sd}))
    xc = x.transpose(2, 0, 1)
    A = oracle.dce_net(xc, oracle.unpack_weights(blob))
    y = oracle.apply_curves(xc, A, sign=-1)
    assert np.abs(y - want[0].numpy()).max() <= 2e-6      # fp32 weights vs the fp64 module
    assert np.abs(A - r[0].numpy()).max() <= 2e-6


def test_state_dict_errors():
    sd = {k: v for k, v in _net().state_dict().items()}
    bad = dict(sd); del bad["e_conv3.bias"]
    with pytest.raises(ValueError, match="e_conv3.bias"):
        L.load_dce_state_dict(bad)
    bad = dict(sd); bad["e_conv7.weight"] = bad["e_conv7.weight"][:, :32]
    with pytest.raises(ValueError, match="e_conv7"):
        L.load_dce_state_dict(bad)
    bad = dict(sd); bad["extra.weight"] = torch.zeros(1)
    with pytest.raises(ValueError, match="unexpected"):
        L.load_dce_state_dict(bad)
    assert L.load_dce_state_dict(bad, strict=False).size == 79416


@pytest.mark.gpu
def test_imported_weights_on_the_gpu():
    net = _net(1)
    x = oracle.synth_lowlight(96, 160, 5)
    with torch.no_grad():
        want, _ = net(torch.from_numpy(x.transpose(2, 0, 1)[None]).double())
    want = want[0].numpy().transpose(1, 2, 0)
    blob = L.load_dce_state_dict(net.state_dict())
    y = L.enhance(x, "zero_dce", dict(weights=blob, curve_sign=-1, precision="fp32"))
    assert np.abs(y - want).max() <= 1e-4
    y = L.enhance(x, "zero_dce", dict(weights=blob, curve_sign=-1, precision="bf16"))
    assert 10 * np.log10(1.0 / np.mean((y - want) ** 2)) >= 50.0
