"""Model zoo structure: variable manifest, shape traces, atrous bookkeeping."""
import pytest
import torch

from tensorflowdistributedlearning_b200 import models as M
from tensorflowdistributedlearning_b200.models.resnet import Block, resnet_v2_beta_block, stack_blocks_plan
from tensorflowdistributedlearning_b200.ops.layers import iter_tf_variables


@pytest.fixture(scope="module")
def seg():
    torch.manual_seed(0)
    return M.ResNetSegModel(model_name="model")


def test_reference_manifest_counts(seg):
    train = [(n, t) for n, t, k in iter_tf_variables(seg) if t.requires_grad]
    stats = [(n, t) for n, t, k in iter_tf_variables(seg) if not t.requires_grad]
    assert len(train) == 208
    assert sum(t.numel() for _, t in train) == 41_814_025
    assert len(stats) == 118
    assert sum(t.numel() for _, t in stats) == 65_632
    assert len({n for n, _ in train}) == 208  # names unique


def test_reference_variable_names(seg):
    names = {n for n, _, _ in iter_tf_variables(seg)}
    for expect in [
        "model/resnet_v2/conv1_1/weights", "model/resnet_v2/conv1_3/BatchNorm/moving_variance",
        "model/resnet_v2/postnorm/gamma", "model/resnet_v2/block1/unit_1/bottleneck_v2/preact/beta",
        "model/resnet_v2/block1/unit_1/bottleneck_v2/shortcut/biases",
        "model/resnet_v2/block2/unit_3/bottleneck_v2/Conv/BatchNorm/beta",
        "model/resnet_v2/block4/unit_2/bottleneck_v2/conv3/weights",
        "model/assp/conv/conv_3x3_2_depthwise/depthwise_weights", "model/assp/conv/conv_3x3_2_depthwise/biases",
        "model/assp/conv/conv_3x3_3_pointwise/BatchNorm/gamma", "model/assp/pooling/conv_1x1/weights",
        "model/assp/conv_1x1/weights", "model/decoder/conv_1x1/BatchNorm/moving_mean",
        "model/decoder/conv_3x3/biases",
    ]:
        assert expect in names, expect


def test_reference_atrous_plan_and_shape_trace(seg):
    assert seg.backbone.unit_plan == [[(1, 1), (1, 1), (2, 1)], [(1, 1)] * 4, [(1, 2)] * 6, [(1, 4), (1, 8), (1, 4)]]
    seg.eval()
    x = torch.randn(1, 101, 101, 2)
    with torch.no_grad():
        _, ep = seg.backbone(x, collect=True)
        y = seg(x)
    assert ep["model/resnet_v2/block1"].shape == (1, 13, 13, 512)
    assert ep["model/resnet_v2/block1/unit_1/bottleneck_v2/conv3"].shape == (1, 26, 26, 512)
    assert ep["model/resnet_v2/block2"].shape == (1, 13, 13, 1032)
    assert ep["model/resnet_v2/block3"].shape == (1, 13, 13, 2048)
    assert ep["model/resnet_v2/block4"].shape == (1, 13, 13, 1024)
    assert y.shape == (1, 101, 101, 1)


def test_stack_blocks_plan_errors():
    blocks = [resnet_v2_beta_block("b1", 8, 2, 2, "bottleneck"), resnet_v2_beta_block("b2", 8, 2, 2, "bottleneck")]
    assert stack_blocks_plan(blocks, None) == [[(1, 1), (2, 1)], [(1, 1), (2, 1)]]
    assert stack_blocks_plan(blocks, 2) == [[(1, 1), (2, 1)], [(1, 1), (1, 1)]]
    with pytest.raises(ValueError):
        stack_blocks_plan(blocks, 8)


def test_strided_unit_equals_conv_then_subsample():
    """Defect 13: the reference convolves at full resolution and then subsamples; the strided conv is the same math."""
    from tensorflowdistributedlearning_b200.ops import functional as TF
    torch.manual_seed(0)
    x = torch.randn(2, 13, 13, 4)
    w = torch.randn(6, 3, 3, 4)
    full = TF.conv2d(x, w, None, 1, 1, "SAME")[:, ::2, ::2, :]
    strided = TF.conv2d(x, w, None, 2, 1, (1, 1, 1, 1))
    assert torch.allclose(full, strided, atol=1e-5)


def test_basic_block_variant_builds_and_runs():
    m = M.ResNetSegModel(model_name="m", n_blocks=(1, 1, 1), block_type="basic_block", base_depth=32)
    m.eval()
    with torch.no_grad():
        y = m(torch.randn(1, 101, 101, 2))
    assert y.shape == (1, 101, 101, 1)


def test_canonical_resnets_param_counts():
    r50 = M.resnet50()
    assert sum(p.numel() for p in r50.parameters()) == 25_557_032
    assert len(list(r50.parameters())) == 161
    r152 = M.resnet152()
    assert sum(p.numel() for p in r152.parameters()) == 60_192_808
    assert len(list(r152.parameters())) == 467


def test_resnet50_forward_shape():
    m = M.resnet50(num_classes=10)
    y = m(torch.randn(2, 64, 64, 3))
    assert y.shape == (2, 10)


def test_xception41_trace_and_params():
    m = M.xception_41(num_classes=1000)
    n = sum(p.numel() for p in m.parameters())
    assert 26.5e6 < n < 27.5e6
    assert len(m.units) == 1 + 1 + 1 + 8 + 1 + 1  # all middle-flow units are built (reference indent bug fixed)
    m.eval()
    with torch.no_grad():
        f = m.features(torch.randn(1, 299, 299, 3))
        y = m(torch.randn(1, 299, 299, 3))
    assert f.shape == (1, 10, 10, 2048)
    assert y.shape == (1, 1000)


def test_xception_output_stride_16():
    m = M.xception_41(num_classes=None, output_stride=16, multi_grid=[1, 2, 4])
    m.eval()
    with torch.no_grad():
        f = m(torch.randn(1, 129, 129, 3))
    assert f.shape[1:3] == (9, 9)


def test_end_points_wanted_fuses_unit_tails_without_changing_values(seg):
    """``wanted`` only changes WHICH tensors are materialised (the others fuse relu(shortcut + conv3) into conv3):
    block outputs, the requested pre-activation end point and the network output stay the same."""
    from tensorflowdistributedlearning_b200.models.resnet import EndPoints

    seg.eval()
    torch.manual_seed(1)
    x = torch.randn(2, 101, 101, 2)
    low = "model/resnet_v2/block1/unit_1/bottleneck_v2/conv3"
    with torch.no_grad():
        net_all, ep_all = seg.backbone(x, collect=True)
        net_w, ep_w = seg.backbone(x, collect=True, wanted=(low,))
        net_none, ep_none = seg.backbone(x, collect=False)
    assert isinstance(ep_w, EndPoints) and ep_none is None
    assert torch.allclose(net_all, net_w, atol=1e-5) and torch.allclose(net_all, net_none, atol=1e-5)
    assert torch.allclose(ep_all[low], ep_w[low], atol=1e-5)
    for k in ("model/resnet_v2/block1", "model/resnet_v2/block4", "model/resnet_v2/block2/unit_3/bottleneck_v2"):
        assert torch.allclose(ep_all[k], ep_w[k], atol=1e-5), k
    # pre-activation tensors nobody asked for are not kept alive
    assert "model/resnet_v2/block2/unit_1/bottleneck_v2/conv3" in ep_all
    assert "model/resnet_v2/block2/unit_1/bottleneck_v2/conv3" not in ep_w


def test_conv2d_residual_relu_fallback_matches_composition():
    from tensorflowdistributedlearning_b200.ops import functional as Fn

    torch.manual_seed(2)
    x = torch.randn(2, 9, 9, 16)
    w = torch.randn(24, 1, 1, 16) / 4
    b = torch.randn(24)
    r = torch.randn(2, 9, 9, 24)
    y = Fn.conv2d(x, w, b, residual=r, relu=True)
    ref = torch.relu(Fn.conv2d(x, w, b) + r)
    assert torch.allclose(y, ref, atol=1e-6)


def test_weight_prep_table_layout_matches_the_c_abi():
    """The batched weight-prep table: 48-byte little-endian records and block prefix sums (include/tdl_b200.h)."""
    import numpy as np

    from tensorflowdistributedlearning_b200.ops import native_ops as O

    w1, w2 = torch.zeros(40, 9, 72), torch.zeros(130, 1, 34)
    f1, f2 = torch.zeros(40, 9, 72, dtype=torch.bfloat16), torch.zeros(130, 1, 34, dtype=torch.bfloat16)
    d1 = torch.zeros(72, 9, 40, dtype=torch.bfloat16)
    descs, starts, total = O.build_weight_prep_table([(w1, f1, d1, 40, 9, 72), (w2, f2, None, 130, 1, 34)])
    assert descs.dtype == torch.uint8 and descs.numel() == 2 * 48
    rec = np.frombuffer(descs.numpy().tobytes(), dtype=np.dtype("<u8,<u8,<u8,<i4,<i4,<i4,<i4,<i4,<i4"))
    assert rec[0][0] == w1.data_ptr() and rec[0][1] == f1.data_ptr() and rec[0][2] == d1.data_ptr()
    assert tuple(rec[0])[3:] == (40, 9, 72, 2, 1, 0)          # 64 x 64 tiles: ceil(72 / 64), ceil(40 / 64)
    assert rec[1][2] == 0 and tuple(rec[1])[3:] == (130, 1, 34, 1, 3, 0)
    assert starts.tolist() == [0, 2 * 1 * 9] and total == 2 * 1 * 9 + 1 * 3 * 1


def test_xception_module_skip_tap_is_only_plumbing(monkeypatch):
    """The skip branch of an Xception module may read an alias handed out by the first depthwise conv (so that its
    gradient is added inside that layer's data-gradient kernel on the GPU).  On the torch path the alias is the input
    itself: outputs and all gradients are identical with the tap on and off, for both skip types."""
    import torch

    from tensorflowdistributedlearning_b200.models.xception import XceptionModule
    from tensorflowdistributedlearning_b200.ops.layers import BNConfig

    for skip, stride, cin in (("sum", 1, 16), ("conv", 2, 8)):
        torch.manual_seed(3)
        m = XceptionModule(cin, [16, 16, 16], skip, stride, [1, 1, 1], 1, False, "u", BNConfig()).train()
        x0 = torch.randn(2, 9, 9, cin)
        res = []
        for tap in ("1", "0"):
            monkeypatch.setenv("TDL_DW_TAP", tap)
            stats = [b.clone() for b in m.buffers()]
            for p in m.parameters():
                p.grad = None
            x = x0.clone().requires_grad_(True)
            y = m(x)
            y.square().sum().backward()
            res.append((y.detach(), x.grad.clone(), [p.grad.clone() for p in m.parameters()]))
            for b, b0 in zip(m.buffers(), stats):
                b.copy_(b0)
        assert torch.equal(res[0][0], res[1][0]) and torch.allclose(res[0][1], res[1][1], atol=1e-6)
        for a, b in zip(res[0][2], res[1][2]):
            assert torch.allclose(a, b, atol=1e-5)


def test_batch_norm_fold_algebra():
    """The identities behind folding a BatchNorm (no activation) into the 1x1 convolution that follows it
    (csrc/bn.cu fold_* kernels): forward conv(scale*y + shift) = conv'(y) + W shift with W' = W * scale; weight
    gradient dW = scale * (dz^T y) + shift x colsum(dz); data gradient w.r.t. the BatchNorm output = dz W (original W)."""
    torch.manual_seed(9)
    y = torch.randn(50, 12, dtype=torch.float64)
    W = torch.randn(7, 12, dtype=torch.float64, requires_grad=True)
    scale = torch.rand(12, dtype=torch.float64) + 0.5
    shift = torch.randn(12, dtype=torch.float64)
    bn_out = (y * scale + shift).requires_grad_(True)
    z = bn_out @ W.t()
    dz = torch.randn_like(z)
    z.backward(dz)
    z_fold = y @ (W.detach() * scale).t() + W.detach() @ shift
    assert torch.allclose(z_fold, z.detach(), atol=1e-10)
    dW_fold = scale * (dz.t() @ y) + shift[None, :] * dz.sum(0)[:, None]
    assert torch.allclose(dW_fold, W.grad, atol=1e-10)
    assert torch.allclose(dz @ W.detach(), bn_out.grad, atol=1e-10)
    # a training-mode BatchNorm behind the convolution makes colsum(dz) vanish and removes W shift from the output
    zc = z.detach() - z.detach().mean(0)
    assert torch.allclose((y @ (W.detach() * scale).t()) - (y @ (W.detach() * scale).t()).mean(0), zc, atol=1e-10)
