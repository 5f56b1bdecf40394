"""ResNet-50 OS16 backbone on the tcgen05 path against the fp32 CPU oracle backbone (torchvision
ResNet-50, what oracle/shim/pytorch_modules.backbones.resnet50 returns) on the same weights and inputs.

Forward features: max|d|/max|ref| <= 5e-2 through 50 bf16 layers (TOL_WHOLE of test_gpu_models.py).  Gradients of a
deep Conv-BN-ReLU stack are bounded against the library bf16 pipeline (the same network under torch autocast on
the GPU): err <= max(2e-2, 1.5 x library error) in L2 norm — see the module docstring of test_gpu_models.py.
"""
import copy

import pytest
import torch
import torchvision

pytestmark = pytest.mark.gpu


def l2_err(got, ref):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    return float((got - ref).norm()) / max(float(ref.norm()), 1e-30)


def rel_err(got, ref):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    return float((got - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)


def _ref_features(net, x):
    f0 = net.relu(net.bn1(net.conv1(x)))
    f1 = net.layer1(net.maxpool(f0))
    f2 = net.layer2(f1)
    f3 = net.layer3(f2)
    f4 = net.layer4(f3)
    return [f0, f1, f2, f3, f4]


@pytest.fixture(scope='module')
def case(lib):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    ref = torchvision.models.resnet50(weights=None, replace_stride_with_dilation=[False, False, True])
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 3, 128, 128, generator=g)
    g1 = torch.randn(2, 256, 32, 32, generator=g)
    g4 = torch.randn(2, 2048, 8, 8, generator=g)
    return ref, x, g1, g4


def test_backbone_train_forward_backward(case):
    from pytorch_segmentation_b200.resnet import ResNet50Features
    ref, x, g1, g4 = case
    ref = copy.deepcopy(ref).train()
    ours = ResNet50Features([False, False, True])
    ours.load_state_dict({k: v for k, v in ref.state_dict().items() if not k.startswith('fc.')})
    ours.cuda().train()
    floor = copy.deepcopy(ref).cuda().train().to(memory_format=torch.channels_last)

    fr = _ref_features(ref, x)
    (fr[1] * g1).sum().add((fr[4] * g4).sum()).backward()

    with torch.autocast('cuda', dtype=torch.bfloat16):
        ff = _ref_features(floor, x.cuda().contiguous(memory_format=torch.channels_last))
    ((ff[1].float() * g1.cuda()).sum() + (ff[4].float() * g4.cuda()).sum()).backward()

    fo = ours(x.cuda())
    assert [tuple(f.shape) for f in fo] == [tuple(f.shape) for f in fr]
    for i in (0, 1, 2, 3, 4):
        e, f = rel_err(fo[i], fr[i]), rel_err(ff[i], fr[i])
        assert e <= max(5e-2, 1.5 * f), 'feature %d: ours %.4f library-bf16 %.4f' % (i, e, f)
    gl = g1.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gh = g4.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    torch.autograd.backward([fo[1], fo[4]], [gl, gh])

    rp, fp, op = dict(ref.named_parameters()), dict(floor.named_parameters()), dict(ours.named_parameters())
    worst = []
    for k, p in op.items():
        assert p.grad is not None, k
        e, f = l2_err(p.grad, rp[k].grad), l2_err(fp[k].grad, rp[k].grad)
        worst.append((e / max(f, 1e-3), k, e, f))
        assert e <= max(2e-2, 1.5 * f), '%s: L2 err ours %.4f vs library-bf16 %.4f' % (k, e, f)
    # running statistics and batch counters follow nn.BatchNorm2d
    rb, ob = dict(ref.named_buffers()), dict(ours.named_buffers())
    for k in ('bn1.running_mean', 'layer1.0.bn1.running_var', 'layer2.0.downsample.1.running_mean',
              'layer4.2.bn3.running_var'):
        assert rel_err(ob[k], rb[k]) < 2e-2, k
    assert int(ob['layer3.1.bn2.num_batches_tracked']) == 1


def test_backbone_eval_forward(case):
    from pytorch_segmentation_b200.resnet import ResNet50Features
    ref, x, _, _ = case
    ref = copy.deepcopy(ref).eval()
    ours = ResNet50Features([False, False, True])
    ours.load_state_dict({k: v for k, v in ref.state_dict().items() if not k.startswith('fc.')})
    ours.cuda().eval()
    with torch.no_grad():
        fr = _ref_features(ref, x)
        fo = ours(x.cuda())
    for i in (1, 4):
        assert rel_err(fo[i], fr[i]) < 5e-2, i
