"""Pix2Pix networks -- drop-in for the hot-path part of the reference's models_pix2pix/networks.py:
define_G / define_D / init_net / init_weights / get_norm_layer / GANLoss / UnetGenerator /
UnetSkipConnectionBlock / NLayerDiscriminator / Cell_upconv / MixedOp_upconv and the module globals
upconv_arch / conv_arch / arch_parameters() (networks.py:23-41,73-281,442-665).

Same constructor signatures, module nesting and state_dict keys (A.4 of SURVEY.md); the torch.nn layers are
parameter containers only -- the arithmetic runs on the sm_100a kernels through gan_program.py.
Out of the hot path (not reachable from the JSRT/ISIC scripts) and therefore not provided: ResnetGenerator,
PixelDiscriminator, cal_gradient_penalty, Cell_conv, the lr schedulers.
"""
import functools

import torch
import torch.nn as nn
from torch.nn import init

from .. import ops
from ..functions import ProgramFunction
from ..gan_program import ARCH_KEY, DiscriminatorProgram, GeneratorProgram

PRIMITIVES_conv = ['conv_421', 'conv_622', 'conv_823']
PRIMITIVES_upconv = ['re_conv_421', 're_conv_622', 're_conv_823']


class Identity(nn.Module):
    def forward(self, x):
        return x


def get_norm_layer(norm_type='instance'):
    if norm_type == 'batch':
        return functools.partial(nn.BatchNorm2d, affine=True, track_running_stats=True)
    if norm_type == 'instance':
        return functools.partial(nn.InstanceNorm2d, affine=False, track_running_stats=False)
    if norm_type == 'none':
        def norm_layer(x):
            return Identity()
        return norm_layer
    raise NotImplementedError('normalization layer [%s] is not found' % norm_type)


def init_weights(net, init_type='normal', init_gain=0.02):
    """Same rule as networks.py:73-104: Conv*/Linear weights by init_type, BatchNorm2d weight N(1, gain)."""
    def init_func(m):
        classname = m.__class__.__name__
        if hasattr(m, 'weight') and (classname.find('Conv') != -1 or classname.find('Linear') != -1):
            if init_type == 'normal':
                init.normal_(m.weight.data, 0.0, init_gain)
            elif init_type == 'xavier':
                init.xavier_normal_(m.weight.data, gain=init_gain)
            elif init_type == 'kaiming':
                init.kaiming_normal_(m.weight.data, a=0, mode='fan_in')
            elif init_type == 'orthogonal':
                init.orthogonal_(m.weight.data, gain=init_gain)
            else:
                raise NotImplementedError('initialization method [%s] is not implemented' % init_type)
            if hasattr(m, 'bias') and m.bias is not None:
                init.constant_(m.bias.data, 0.0)
        elif classname.find('BatchNorm2d') != -1:
            init.normal_(m.weight.data, 1.0, init_gain)
            init.constant_(m.bias.data, 0.0)

    print('initialize network with %s' % init_type)
    net.apply(init_func)


def init_net(net, init_type='normal', init_gain=0.02, gpu_ids=[]):
    """networks.py:107-122.  With gpu_ids the net moves to gpu_ids[0] and is wrapped in nn.DataParallel exactly
    like the reference (single id = pass-through; checkpoints keep their 'module.' prefix)."""
    if len(gpu_ids) > 0:
        assert (torch.cuda.is_available())
        net.to(gpu_ids[0])
        net = torch.nn.DataParallel(net, gpu_ids)
    init_weights(net, init_type, init_gain=init_gain)
    return net


def define_G(input_nc, output_nc, ngf, netG, norm='batch', use_dropout=False, init_type='normal', init_gain=0.02,
             gpu_ids=[]):
    norm_layer = get_norm_layer(norm_type=norm)
    if netG == 'unet_256':
        net = UnetGenerator(input_nc, output_nc, 8, ngf, norm_layer=norm_layer, use_dropout=use_dropout)
    elif netG == 'unet_128':
        net = UnetGenerator(input_nc, output_nc, 7, ngf, norm_layer=norm_layer, use_dropout=use_dropout)
    elif netG in ('resnet_9blocks', 'resnet_6blocks'):
        raise NotImplementedError('genseg_b200: generator [%s] is outside the B200 hot path (pix2pix_model.py:36 '
                                  'sets unet_256)' % netG)
    else:
        raise NotImplementedError('Generator model name [%s] is not recognized' % netG)
    return init_net(net, init_type, init_gain, gpu_ids)


def define_D(input_nc, ndf, netD, n_layers_D=3, norm='batch', init_type='normal', init_gain=0.02, gpu_ids=[]):
    norm_layer = get_norm_layer(norm_type=norm)
    if netD == 'basic':
        net = NLayerDiscriminator(input_nc, ndf, n_layers=3, norm_layer=norm_layer)
    elif netD == 'n_layers':
        net = NLayerDiscriminator(input_nc, ndf, n_layers_D, norm_layer=norm_layer)
    elif netD == 'pixel':
        raise NotImplementedError('genseg_b200: PixelDiscriminator is outside the B200 hot path')
    else:
        raise NotImplementedError('Discriminator model name [%s] is not recognized' % netD)
    return init_net(net, init_type, init_gain, gpu_ids)


# ---------------------------------------------------------------------------------------------------------
# losses
# ---------------------------------------------------------------------------------------------------------
class _GanLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, mode, label):
        p = pred.contiguous()
        if p.dtype != torch.float32:
            p = p.float()
        ctx.save_for_backward(p)
        ctx.meta = (mode, label)
        return ops.gan_loss_fwd(p, mode, label)

    @staticmethod
    def backward(ctx, g):
        (p,) = ctx.saved_tensors
        mode, label = ctx.meta
        return ops.gan_loss_bwd(p, mode, label, g.contiguous().float()), None, None


class _L1LossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        a, b = a.contiguous().float(), b.contiguous().float()
        ctx.save_for_backward(a, b)
        return ops.l1_loss_fwd(a, b)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        d = ops.l1_loss_bwd(a, b, g.contiguous().float())
        return d, (-d if ctx.needs_input_grad[1] else None)


def l1_loss(input, target):
    """mean |input - target| (torch.nn.L1Loss of pix2pix_model.py:73,127) on the fused kernel."""
    if not input.is_cuda:
        raise RuntimeError("genseg_b200: l1_loss needs CUDA tensors; there is no CPU fallback")
    return _L1LossFn.apply(input, target)


class GANLoss(nn.Module):
    """networks.py:215-281: vanilla (BCE-with-logits), lsgan (MSE) or wgangp (+-mean) against constant labels."""

    def __init__(self, gan_mode, target_real_label=1.0, target_fake_label=0.0):
        super(GANLoss, self).__init__()
        self.register_buffer('real_label', torch.tensor(target_real_label))
        self.register_buffer('fake_label', torch.tensor(target_fake_label))
        self.gan_mode = gan_mode
        self._real, self._fake = float(target_real_label), float(target_fake_label)
        if gan_mode not in ('lsgan', 'vanilla', 'wgangp'):
            raise NotImplementedError('gan mode %s not implemented' % gan_mode)
        self.loss = None

    def get_target_tensor(self, prediction, target_is_real):
        target_tensor = self.real_label if target_is_real else self.fake_label
        return target_tensor.expand_as(prediction)

    def __call__(self, prediction, target_is_real):
        if not prediction.is_cuda:
            raise RuntimeError("genseg_b200: GANLoss needs CUDA tensors; there is no CPU fallback")
        if self.gan_mode == 'wgangp':
            return _GanLossFn.apply(prediction, 2, 1.0 if target_is_real else 0.0)
        label = self._real if target_is_real else self._fake
        return _GanLossFn.apply(prediction, 0 if self.gan_mode == 'vanilla' else 1, label)


# ---------------------------------------------------------------------------------------------------------
# architecture parameters (module globals, networks.py:442-484)
# ---------------------------------------------------------------------------------------------------------
num_ops_conv = len(PRIMITIVES_conv)
conv_arch = (1e-3 * torch.randn(8, num_ops_conv)).requires_grad_(True)
_conv_arch_parameters = [conv_arch, ]


def conv_arch_parameters():
    return _conv_arch_parameters


num_ops_upconv = len(PRIMITIVES_upconv)
upconv_arch = (1e-3 * torch.randn(8, num_ops_upconv)).requires_grad_(True)
_upconv_arch_parameters = [upconv_arch, ]


def upconv_arch_parameters():
    return _upconv_arch_parameters


_arch_parameters = [upconv_arch, conv_arch]


def arch_parameters():
    return _arch_parameters


class _ReConv(nn.Module):
    """architecture_pix2pix/operations.py:14-39 (re_conv_421 / 622 / 823): container of one ConvTranspose2d."""

    def __init__(self, C_in, C_out, bias, k):
        super().__init__()
        self.op = nn.ConvTranspose2d(in_channels=C_in, out_channels=C_out, kernel_size=k, stride=2, padding=(k - 2) // 2,
                                     bias=bias)


class re_conv_421(_ReConv):
    def __init__(self, C_in, C_out, bias):
        super().__init__(C_in, C_out, bias, 4)


class re_conv_622(_ReConv):
    def __init__(self, C_in, C_out, bias):
        super().__init__(C_in, C_out, bias, 6)


class re_conv_823(_ReConv):
    def __init__(self, C_in, C_out, bias):
        super().__init__(C_in, C_out, bias, 8)


OPS = {'re_conv_421': re_conv_421, 're_conv_622': re_conv_622, 're_conv_823': re_conv_823}


class MixedOp_upconv(nn.Module):
    def __init__(self, C_in, C_out, bias):
        super(MixedOp_upconv, self).__init__()
        self._ops = nn.ModuleList()
        for primitive in PRIMITIVES_upconv:
            self._ops.append(OPS[primitive](C_in, C_out, bias))


class Cell_upconv(nn.Module):
    def __init__(self, C_in, C_out, bias, layer_index):
        super(Cell_upconv, self).__init__()
        self._layer_index = layer_index
        self._ops = MixedOp_upconv(C_in, C_out, bias)


class UnetSkipConnectionBlock(nn.Module):
    """networks.py:553-617 -- structure/parameter container (the whole generator runs as one program)."""

    def __init__(self, outer_nc, inner_nc, input_nc=None, layer_index=None, submodule=None, outermost=False,
                 innermost=False, norm_layer=nn.BatchNorm2d, use_dropout=False):
        super(UnetSkipConnectionBlock, self).__init__()
        self.outermost = outermost
        if type(norm_layer) == functools.partial:
            use_bias = norm_layer.func == nn.InstanceNorm2d
        else:
            use_bias = norm_layer == nn.InstanceNorm2d
        if input_nc is None:
            input_nc = outer_nc
        downconv = nn.Conv2d(input_nc, inner_nc, kernel_size=4, stride=2, padding=1, bias=use_bias)
        downrelu = nn.LeakyReLU(0.2, True)
        downnorm = norm_layer(inner_nc)
        uprelu = nn.ReLU(True)
        upnorm = norm_layer(outer_nc)
        if outermost:
            upconv = Cell_upconv(inner_nc * 2, outer_nc, bias=True, layer_index=layer_index)
            model = [downconv] + [submodule] + [uprelu, upconv, nn.Tanh()]
        elif innermost:
            upconv = Cell_upconv(inner_nc, outer_nc, bias=use_bias, layer_index=layer_index)
            model = [downrelu, downconv] + [uprelu, upconv, upnorm]
        else:
            upconv = Cell_upconv(inner_nc * 2, outer_nc, bias=use_bias, layer_index=layer_index)
            model = [downrelu, downconv, downnorm] + [submodule] + [uprelu, upconv, upnorm]
            if use_dropout:
                model = model + [nn.Dropout(0.5)]
        self.model = nn.Sequential(*model)


class UnetGenerator(nn.Module):
    """Create a Unet-based generator (networks.py:514-550); forward = one program on the sm_100a kernels."""

    def __init__(self, input_nc, output_nc, num_downs, ngf=64, norm_layer=nn.BatchNorm2d, use_dropout=False):
        super(UnetGenerator, self).__init__()
        self.layer_index = 0
        unet_block = UnetSkipConnectionBlock(ngf * 8, ngf * 8, input_nc=None, layer_index=self.layer_index, submodule=None,
                                             norm_layer=norm_layer, innermost=True)
        self.layer_index += 1
        for i in range(num_downs - 5):
            unet_block = UnetSkipConnectionBlock(ngf * 8, ngf * 8, input_nc=None, layer_index=self.layer_index,
                                                 submodule=unet_block, norm_layer=norm_layer, use_dropout=use_dropout)
            self.layer_index += 1
        for mult_o, mult_i in ((4, 8), (2, 4), (1, 2)):
            unet_block = UnetSkipConnectionBlock(ngf * mult_o, ngf * mult_i, input_nc=None, layer_index=self.layer_index,
                                                 submodule=unet_block, norm_layer=norm_layer)
            self.layer_index += 1
        self.model = UnetSkipConnectionBlock(output_nc, ngf, input_nc=input_nc, layer_index=self.layer_index,
                                             submodule=unet_block, outermost=True, norm_layer=norm_layer)
        self.layer_index += 1
        inst = (norm_layer.func if type(norm_layer) == functools.partial else norm_layer) == nn.InstanceNorm2d
        self._program = GeneratorProgram(input_nc, output_nc, ngf, use_dropout, instance_norm=inst, num_downs=num_downs)
        self._dropout_masks = None      # test hook: list of three uint8 NHWC keep-masks (depth 4, 5, 6)

    def forward(self, input):
        if not input.is_cuda:
            raise RuntimeError("genseg_b200: the B200 hot path needs CUDA tensors; there is no CPU fallback")
        import sys
        arch = sys.modules[__name__].upconv_arch      # callers rebind the module global (pix2pix_model.py:63-64)
        if arch.device != input.device:
            raise RuntimeError("networks.upconv_arch lives on %s but the input is on %s (the reference moves it with "
                               "networks.upconv_arch = networks.upconv_arch.cuda())" % (arch.device, input.device))
        pd = dict(self.named_parameters())
        bd = dict(self.named_buffers())
        params = [pd[k] for k in self._program.param_names if k != ARCH_KEY] + [arch]
        extra = {"drop_masks": self._dropout_masks} if self._dropout_masks is not None else None
        return ProgramFunction.apply(self._program, bd, self.training, extra, 1, torch.is_grad_enabled(), input, *params)


class NLayerDiscriminator(nn.Module):
    """Defines a PatchGAN discriminator (networks.py:620-665)."""

    def __init__(self, input_nc, ndf=64, n_layers=3, norm_layer=nn.BatchNorm2d):
        super(NLayerDiscriminator, self).__init__()
        if type(norm_layer) == functools.partial:
            use_bias = norm_layer.func == nn.InstanceNorm2d
        else:
            use_bias = norm_layer == nn.InstanceNorm2d
        kw, padw = 4, 1
        sequence = [nn.Conv2d(input_nc, ndf, kernel_size=kw, stride=2, padding=padw), nn.LeakyReLU(0.2, True)]
        nf_mult = 1
        for n in range(1, n_layers):
            nf_mult_prev, nf_mult = nf_mult, min(2 ** n, 8)
            sequence += [nn.Conv2d(ndf * nf_mult_prev, ndf * nf_mult, kernel_size=kw, stride=2, padding=padw, bias=use_bias),
                         norm_layer(ndf * nf_mult), nn.LeakyReLU(0.2, True)]
        nf_mult_prev, nf_mult = nf_mult, min(2 ** n_layers, 8)
        sequence += [nn.Conv2d(ndf * nf_mult_prev, ndf * nf_mult, kernel_size=kw, stride=1, padding=padw, bias=use_bias),
                     norm_layer(ndf * nf_mult), nn.LeakyReLU(0.2, True)]
        sequence += [nn.Conv2d(ndf * nf_mult, 1, kernel_size=kw, stride=1, padding=padw)]
        self.model = nn.Sequential(*sequence)
        self._program = DiscriminatorProgram(input_nc, ndf, n_layers, instance_norm=use_bias)

    def forward(self, input):
        from ..functions import run_program
        return run_program(self._program, self, input)
