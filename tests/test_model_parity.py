"""Model parity against the reference model used as a CPU oracle."""
import torch

from cs744_ddp_b200.models import vgg


def test_init_parity_all_58_tensors(ref_model_module):
    torch.manual_seed(0)
    ref = ref_model_module.VGG11()
    torch.manual_seed(0)
    ours = vgg.VGG11()
    rs, os_ = ref.state_dict(), ours.state_dict()
    assert list(rs.keys()) == list(os_.keys())
    assert len(os_) == 58
    for k in rs:
        assert rs[k].dtype == os_[k].dtype and rs[k].shape == os_[k].shape, k
        assert torch.equal(rs[k], os_[k]), k


def test_state_dict_key_table_matches_module():
    for name in vgg.VGG_PLANS:
        m = vgg.VGG(name)
        assert list(m.state_dict().keys()) == vgg.state_dict_keys(name)
        assert [n for n, _ in m.named_parameters()] == [s.name for s in vgg.param_specs(name)]
        for s, (_, p) in zip(vgg.param_specs(name), m.named_parameters()):
            assert tuple(p.shape) == s.shape


def test_vgg11_parameter_count():
    specs = vgg.param_specs("VGG11")
    assert len(specs) == 34
    assert sum(s.numel for s in specs) == 9_231_114


def test_forward_backward_parity(ref_model_module):
    torch.manual_seed(0)
    ref = ref_model_module.VGG11()
    torch.manual_seed(0)
    ours = vgg.VGG11()
    x = torch.randn(8, 3, 32, 32)
    y = torch.randint(0, 10, (8,))
    lr_ = torch.nn.functional.cross_entropy(ref(x), y)
    lo = torch.nn.functional.cross_entropy(ours(x), y)
    assert torch.equal(lr_, lo)
    lr_.backward()
    lo.backward()
    for (n, a), (_, b) in zip(ref.named_parameters(), ours.named_parameters()):
        assert torch.equal(a.grad, b.grad), n


def test_checkpoint_roundtrip_with_reference(ref_model_module, tmp_path):
    from cs744_ddp_b200.utils import checkpoint as ckpt
    ours = vgg.VGG11()
    path = tmp_path / "m.pt"
    ckpt.save_model(str(path), ours.state_dict())
    ref = ref_model_module.VGG11()
    ref.load_state_dict(torch.load(path), strict=True)
    # a DDP-wrapped reference model writes "module."-prefixed keys: we accept those
    ckpt.save_model(str(path), ref.state_dict(), ddp_prefix=True)
    blob = ckpt.load(str(path))
    ours2 = vgg.VGG11()
    ours2.load_state_dict(blob["model"], strict=True)
    for k, v in ref.state_dict().items():
        assert torch.equal(v, ours2.state_dict()[k])


def test_grad_ready_order_matches_autograd():
    m = vgg.VGG11()
    seen = []
    for n, p in m.named_parameters():
        p.register_post_accumulate_grad_hook(lambda _p, n=n: seen.append(n))
    m(torch.randn(2, 3, 32, 32)).sum().backward()
    assert seen == vgg.grad_ready_order("VGG11")


def test_layout_roundtrip():
    w = torch.randn(8, 3, 3, 3)
    assert torch.equal(vgg.krsc_to_oihw(vgg.oihw_to_krsc(w)), w)


def test_flops():
    assert abs(vgg.forward_flops_per_image("VGG11") / 1e6 - 305.5) < 0.5
