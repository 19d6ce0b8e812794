"""Model-zoo parity (SURVEY §7.4 'Model parity'): keys, shapes, dtypes, order and CPU logits
against the reference's own modules."""
import pytest
import torch

from fedb200 import models

# name in our registry -> expression on the reference package
REF_CTORS = {
    "LeNet": "LeNet()", "MobileNet": "MobileNet()", "MobileNetV2": "MobileNetV2()",
    "ResNet18": "ResNet18()", "ResNet34": "ResNet34()", "ResNet50": "ResNet50()",
    "ResNet101": "ResNet101()", "ResNet152": "ResNet152()",
    "PreActResNet18": "PreActResNet18()", "PreActResNet34": "PreActResNet34()",
    "PreActResNet50": "PreActResNet50()", "PreActResNet101": "PreActResNet101()",
    "PreActResNet152": "PreActResNet152()",
    "VGG11": "VGG('VGG11')", "VGG13": "VGG('VGG13')", "VGG16": "VGG('VGG16')", "VGG19": "VGG('VGG19')",
    "GoogLeNet": "GoogLeNet()",
    "DenseNet121": "DenseNet121()", "DenseNet169": "DenseNet169()", "DenseNet201": "DenseNet201()",
    "DenseNet161": "DenseNet161()", "densenet_cifar": "densenet_cifar()",
    "ResNeXt29_2x64d": "ResNeXt29_2x64d()", "ResNeXt29_4x64d": "ResNeXt29_4x64d()",
    "ResNeXt29_8x64d": "ResNeXt29_8x64d()", "ResNeXt29_32x4d": "ResNeXt29_32x4d()",
    "DPN26": "DPN26()", "DPN92": "DPN92()", "SENet18": "SENet18()",
    "ShuffleNetV2_0.5": "ShuffleNetV2(0.5)", "ShuffleNetV2_1": "ShuffleNetV2(1)",
    "ShuffleNetV2_1.5": "ShuffleNetV2(1.5)", "ShuffleNetV2_2": "ShuffleNetV2(2)",
    "EfficientNetB0": "EfficientNetB0()",
    "RegNetX_200MF": "RegNetX_200MF()", "RegNetX_400MF": "RegNetX_400MF()",
    "RegNetY_400MF": "RegNetY_400MF()",
    "PNASNetA": "PNASNetA()", "PNASNetB": "PNASNetB()", "SimpleDLA": "SimpleDLA()", "DLA": "DLA()",
}
# big ones: layout check only (keeps the CPU suite fast)
HEAVY = {"ResNet101", "ResNet152", "PreActResNet101", "PreActResNet152", "DenseNet169", "DenseNet201",
         "DenseNet161", "ResNeXt29_4x64d", "ResNeXt29_8x64d", "DPN92", "ResNet50", "PreActResNet50",
         "ResNet34", "PreActResNet34", "VGG16", "VGG19", "DenseNet121", "GoogLeNet", "DLA",
         "SimpleDLA", "DPN26", "ResNeXt29_2x64d"}


def test_registry_covers_all_44():
    assert len(models.REGISTRY) == 44
    assert set(REF_CTORS) | {"ShuffleNetG2", "ShuffleNetG3"} == set(models.REGISTRY)


@pytest.mark.parametrize("name", sorted(REF_CTORS))
def test_state_dict_layout_matches_reference(name, ref_models):
    torch.manual_seed(0)
    ref = eval("ref_models." + REF_CTORS[name], {"ref_models": ref_models})
    torch.manual_seed(0)
    mine = models.build(name)
    a, b = ref.state_dict(), mine.state_dict()
    assert list(a.keys()) == list(b.keys())
    for k in a:
        assert a[k].shape == b[k].shape and a[k].dtype == b[k].dtype, k
    # same construction order => same RNG stream => identical initial weights
    for k in a:
        assert torch.equal(a[k], b[k]), k
    if name in HEAVY:
        return
    x = torch.randn(2, 3, 32, 32)
    ref.eval(), mine.eval()
    with torch.no_grad():
        torch.testing.assert_close(mine(x), ref(x), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("name", ["LeNet", "MobileNet", "ResNet18", "SENet18", "EfficientNetB0",
                                  "ShuffleNetV2_0.5", "PNASNetA", "RegNetY_400MF", "densenet_cifar"])
def test_train_step_matches_reference(name, ref_models):
    """fwd + bwd + BN running stats in train mode, CPU."""
    torch.manual_seed(1)
    ref = eval("ref_models." + REF_CTORS[name], {"ref_models": ref_models})
    mine = models.build(name)
    mine.load_state_dict(ref.state_dict())
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
    ref.train(), mine.train()
    if name == "EfficientNetB0":  # dropout draws: align the generators
        torch.manual_seed(5)
    lr = torch.nn.functional.cross_entropy(ref(x), y)
    if name == "EfficientNetB0":
        torch.manual_seed(5)
    lm = torch.nn.functional.cross_entropy(mine(x), y)
    lr.backward(), lm.backward()
    torch.testing.assert_close(lm, lr, rtol=1e-4, atol=1e-5)
    rp, mp = dict(ref.named_parameters()), dict(mine.named_parameters())
    for k, p in rp.items():
        if p.grad is None:
            assert mp[k].grad is None, k
            continue
        torch.testing.assert_close(mp[k].grad, p.grad, rtol=2e-3, atol=2e-5, msg=k)
    for (k, a), (_, b) in zip(ref.state_dict().items(), mine.state_dict().items()):
        torch.testing.assert_close(b, a, rtol=1e-4, atol=1e-5, msg=k)


@pytest.mark.parametrize("name", ["ShuffleNetG2", "ShuffleNetG3"])
def test_shufflenet_g_constructs_and_runs(name):
    """broken upstream, fixed here with integer division."""
    net = models.build(name).eval()
    with torch.no_grad():
        assert net(torch.randn(2, 3, 32, 32)).shape == (2, 10)
