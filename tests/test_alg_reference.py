"""The algebra behind csrc/conv_in_alg.cu on the CPU, in fp64: the first layer's weight gradient written with the input patches'
Gram matrix and one patch per pooling window (ops/reference.py::conv1_wgrad_from_pooled) equals autograd through
Conv3d -> BatchNorm3d -> ReLU -> MaxPool3d.  (The kernels themselves are checked on the GPU: tests/test_alg_bwd0_gpu.py.)"""
import pytest
import torch
import torch.nn.functional as F

from neuroimagedisttraining_b200.ops.reference import conv1_wgrad_from_pooled


@pytest.mark.parametrize("shape,bias", [((17, 19, 21), True), ((23, 17, 17), False)])
def test_gram_form_of_the_first_layer_gradient_equals_autograd(shape, bias):
    torch.manual_seed(0)
    N, C = 3, 8
    x = torch.rand(N, 1, *shape, dtype=torch.float64)
    x[:, :, :, :, :4] = 0.0                                         # background
    w = (torch.randn(C, 1, 5, 5, 5, dtype=torch.float64) * 0.1).requires_grad_(True)
    b = (torch.randn(C, dtype=torch.float64) * 0.1).requires_grad_(True) if bias else None
    gamma = torch.rand(C, dtype=torch.float64) + 0.5
    gamma[2] = -0.7                                                 # the pooling winner is then the window's minimum of y
    gamma = gamma.requires_grad_(True)
    beta = (torch.randn(C, dtype=torch.float64) * 0.2).requires_grad_(True)
    out = F.max_pool3d(F.relu(F.batch_norm(F.conv3d(x, w, b, stride=2), None, None, gamma, beta, training=True, eps=1e-5)), 3, 3)
    da = torch.randn_like(out)
    out.backward(da)
    dW, dgamma = conv1_wgrad_from_pooled(x, w.detach(), None if b is None else b.detach(), gamma.detach(), beta.detach(), da)
    torch.testing.assert_close(dW, w.grad, rtol=1e-8, atol=1e-10)
    torch.testing.assert_close(dgamma, gamma.grad, rtol=1e-8, atol=1e-10)
    if b is not None:                                               # BatchNorm removes the batch mean: the conv bias gets no gradient
        assert float(b.grad.abs().max()) < 1e-10
