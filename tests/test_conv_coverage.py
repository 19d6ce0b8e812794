"""Shape logic of the convolution lowering (CPU): which kernel family every zoo convolution is routed to at batch 128.
Guards the tensor-core coverage reached in round 1 against regressions of the predicates in fedb200/ops/cuda_gemm.py."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import conv_coverage as cc  # noqa: E402


def _shape(N, H, C, K, R, stride, groups, pad=None):
    pad = R // 2 if pad is None else pad
    P = (H + 2 * pad - R) // stride + 1
    return (N, H, H, C, K, R, R, P, P, stride, stride, pad, pad, groups)


@pytest.mark.parametrize("shape,want", [
    (_shape(128, 32, 64, 64, 3, 1, 1), ("tc", "tc", "tc")),                    # ResNet layer1
    (_shape(128, 32, 64, 128, 3, 2, 1), ("tc", "tc", "tc")),                   # down-sampling 3x3
    (_shape(128, 32, 64, 128, 1, 2, 1), ("tc", "tc", "tc")),                   # projection shortcut
    (_shape(128, 32, 3, 64, 3, 1, 1), ("tc-im2col", "-", "tc-im2col")),        # stem
    (_shape(128, 32, 3, 44, 3, 1, 1), ("tc-im2col", "-", "tc-im2col")),        # PNASNet stem: Cout not a multiple of 64
    (_shape(128, 32, 16, 96, 1, 1, 1), ("tc", "tc", "tc")),                    # few input channels: a GEMM, not im2col
    (_shape(128, 32, 48, 12, 3, 1, 1), ("tc", "tc", "tc")),                    # DenseNet growth 12: padded bf16 rows
    (_shape(128, 16, 116, 116, 1, 1, 1), ("tc", "tc", "tc")),                  # ShuffleNetV2 x1
    (_shape(128, 32, 32, 32, 3, 1, 32), ("depthwise",) * 3),
    (_shape(128, 16, 96, 96, 5, 2, 96), ("depthwise",) * 3),                   # EfficientNet 5x5
    (_shape(128, 8, 88, 88, 7, 1, 88), ("depthwise",) * 3),                    # PNASNet 7x7
    (_shape(128, 32, 128, 128, 3, 1, 2), ("tc-grouped",) * 3),                 # ResNeXt 2x64d
    (_shape(128, 8, 152, 152, 3, 2, 19), ("tc-grouped",) * 3),                 # RegNetX_200MF, 19 groups of 8
    (_shape(128, 32, 128, 128, 3, 1, 32), ("tc-grouped", "tc-grouped", "generic")),   # 4-wide groups: bf16 pitch too small
    (_shape(128, 16, 58, 58, 1, 1, 1), ("generic", "generic", "tc")),          # 58 channels: no 16-byte fp32 pitch
    (_shape(128, 32, 6, 16, 5, 1, 1, pad=0), ("generic",) * 3),                # 5x5 dense (LeNet has its own kernels)
])
def test_lowering_of_representative_layers(shape, want):
    assert cc.classify(shape) == want


@pytest.mark.parametrize("name,limit", [("ResNet18", 0.0), ("VGG16", 0.0), ("MobileNet", 0.0), ("DenseNet121", 0.0),
                                        ("GoogLeNet", 0.0), ("EfficientNetB0", 0.0), ("ResNeXt29_2x64d", 0.0),
                                        ("RegNetX_200MF", 0.0), ("densenet_cifar", 0.0), ("PNASNetA", 0.07), ("DPN26", 0.03)])
def test_share_of_flops_left_on_the_generic_kernels(name, limit):
    shares = cc.generic_share(name)
    assert max(shares.values()) <= limit + 1e-3, shares      # 0.1 %: the tiny squeeze-excite 1x1 layers on 1x1 maps
