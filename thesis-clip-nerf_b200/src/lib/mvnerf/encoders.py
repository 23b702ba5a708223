"""Upstream producers of ``combined_features``: torch mirrors of the reference's
``VisualFeatures`` (``src/lib/mvnerf/layers.py:7-259``) and ``CombineCLIPVisualV4`` / ``V3`` (``:414-660``).

These run once per scene, before the per-ray hot path, and are ordinary convolutions / attention: they are built from
torch modules (cuDNN / cuBLAS / the library attention kernel) and are NOT part of the hand-written CUDA path.  They
exist so that the drop-in models are complete -- ``MVVNeRFRenderer.visual_features`` / ``combine_clip_visual`` are real
modules with the attribute paths ``src/train_nerf.py:27-32`` builds its optimizer groups from -- and so that the gather's
low-resolution mode has its producer: ``CombineCLIPVisual*(final_upsample=False)`` returns the 240 x 320 input of the
reference's last ``UpSampling2D`` (``layers.py:589,659``), which ``CnScene.feat_up = 2`` consumes.

Interface = the reference's: channels-LAST tensors in and out (``[(b n), H, W, C]``), the same constructor keywords and
sub-module attribute names.  Keras semantics that differ from torch defaults are written out:
  * ``padding='same'`` with a stride pads asymmetrically (the extra row / column at the end);
  * ``BatchNormalization`` eps 1e-3, momentum 0.99; ``Block`` always normalises with batch statistics
    (``training=True`` is hard-wired, ``layers.py:22,26``) and uses ONE BatchNormalization for both convolutions
    (``norm_1`` is assigned twice, ``:11,14``); ``TransformerBlock.layer_norm_1`` is a BatchNormalization too (``:76``);
  * ``LayerNormalization`` eps 1e-3; Dense ``activation='gelu'`` is the exact (erf) form;
  * ``Resizing`` / ``UpSampling2D(interpolation='bilinear')`` sample at half-pixel centres without antialiasing
    (= ``F.interpolate(mode='bilinear', align_corners=False)``; proven against the oracle's restatement in the tests);
  * ``MultiHeadAttention(num_heads, key_dim)``: per-head projections with bias, scores scaled by key_dim ** -0.5.
The CLIP encoders themselves are a frozen TensorFlow SavedModel at an absolute path outside the reference repository
(``src/lib/clip/model.py:5-8``): they stay a plug-in (``model.clip_visual`` = any callable returning the 5-tuple
``(embedding [B,1024], [B,56,56,256], [B,28,28,512], [B,14,14,1024], [B,7,7,2048])``).

Checkpoints: ``keras_variables(module)`` / ``load_keras_variables(module, mapping)`` exchange every variable in Keras
layout (conv kernels HWIO, Dense ``[in, out]``, attention kernels ``[embed, heads, dim]``) under the reference's
attribute paths; ``tools/convert_tf_checkpoint.py`` (run where TensorFlow is installed) writes such a mapping from a
reference checkpoint as ``.npz``."""
from __future__ import annotations

import math
from typing import Dict, Sequence

import torch
import torch.nn.functional as F

_BN_EPS, _BN_MOMENTUM, _LN_EPS = 1e-3, 0.01, 1e-3        # torch momentum = 1 - keras momentum (0.99)


def _nchw(x):
    return x.permute(0, 3, 1, 2)


def _nhwc(x):
    return x.permute(0, 2, 3, 1)


def same_pad(x: torch.Tensor, kernel: int, stride: int) -> torch.Tensor:
    """TensorFlow ``padding='same'`` for NCHW ``x``: total = max((ceil(n / s) - 1) s + k - n, 0), the odd one at the end"""
    pads = []
    for n in (x.shape[-1], x.shape[-2]):                  # F.pad takes the last dimension first
        total = max((math.ceil(n / stride) - 1) * stride + kernel - n, 0)
        pads += [total // 2, total - total // 2]
    return F.pad(x, pads) if any(pads) else x


def resize_bilinear(x: torch.Tensor, size) -> torch.Tensor:
    """``tf.keras.layers.Resizing(h, w, 'bilinear')`` / ``UpSampling2D(interpolation='bilinear')`` on NCHW"""
    return F.interpolate(x, size=tuple(size), mode='bilinear', align_corners=False, antialias=False)


class Conv2D(torch.nn.Conv2d):
    """Keras ``Conv2D(filters, kernel_size, strides, padding)`` on NCHW tensors (glorot_uniform kernel, zero bias)"""

    def __init__(self, in_ch, filters, kernel_size, strides=1, padding='valid', use_bias=True, activation=None):
        super().__init__(in_ch, filters, kernel_size, stride=strides, padding=0, bias=use_bias)
        self.same = padding == 'same'
        self.activation = activation
        torch.nn.init.xavier_uniform_(self.weight)
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)

    def forward(self, x):
        if self.same:
            x = same_pad(x, self.kernel_size[0], self.stride[0])
        y = super().forward(x)
        return _activate(y, self.activation)


def _activate(x, name):
    if name is None:
        return x
    if name == 'relu':
        return torch.relu(x)
    if name == 'elu':
        return F.elu(x)
    raise ValueError(f'activation {name} not supported')


class BatchNorm(torch.nn.Module):
    """Keras ``BatchNormalization`` over the channel axis of an NCHW / NCL tensor; ``always_batch_stats`` reproduces
    a call with ``training=True`` hard-wired"""

    def __init__(self, channels, always_batch_stats=False):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(channels))
        self.bias = torch.nn.Parameter(torch.zeros(channels))
        self.register_buffer('running_mean', torch.zeros(channels))
        self.register_buffer('running_var', torch.ones(channels))
        self.always = always_batch_stats

    def forward(self, x):
        return F.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias,
                            training=self.always or self.training, momentum=_BN_MOMENTUM, eps=_BN_EPS)


class Block(torch.nn.Module):
    """``layers.py:7-34``: conv-norm-relu-conv-norm + skip, relu; ONE shared BatchNormalization, batch statistics always"""

    def __init__(self, in_ch, n_features, downsample=None):
        super().__init__()
        self.conv_1 = Conv2D(in_ch, n_features, 3, padding='same')
        self.conv_2 = Conv2D(n_features, n_features, 3, padding='same')
        self.norm_1 = BatchNorm(n_features, always_batch_stats=True)
        self.downsample = downsample

    def forward(self, x):
        out = torch.relu(self.norm_1(self.conv_1(x)))
        out = self.norm_1(self.conv_2(out))
        skip = x if self.downsample is None else self.downsample(x)
        return torch.relu(out + skip)


class ConvolutionalEncoder(torch.nn.Module):
    """``layers.py:37-59``: 7x7 stride-2 stem to 64 channels, three residual blocks of n_features / 2 -> [B, H/2, W/2, 128]"""

    def __init__(self, n_features):
        super().__init__()
        h = n_features // 2
        downsample = torch.nn.Sequential(Conv2D(64, h, 1, padding='same', use_bias=False), BatchNorm(h))
        self.conv_features = torch.nn.Sequential(
            Conv2D(3, 64, 7, strides=2, padding='same', use_bias=False), BatchNorm(64), torch.nn.ReLU(),
            Block(64, h, downsample), Block(h, h), Block(h, h))

    def forward(self, images_nhwc):
        return _nhwc(self.conv_features(_nchw(images_nhwc)))


class MultiHeadAttention(torch.nn.Module):
    """Keras ``MultiHeadAttention(num_heads, key_dim, value_dim)`` self-attention on [B, T, E]"""

    def __init__(self, embed_dim, num_heads, key_dim):
        super().__init__()
        self.h, self.d = num_heads, key_dim
        self.query = torch.nn.Linear(embed_dim, num_heads * key_dim)
        self.key = torch.nn.Linear(embed_dim, num_heads * key_dim)
        self.value = torch.nn.Linear(embed_dim, num_heads * key_dim)
        self.output = torch.nn.Linear(num_heads * key_dim, embed_dim)
        for m in (self.query, self.key, self.value, self.output):
            torch.nn.init.xavier_uniform_(m.weight)
            torch.nn.init.zeros_(m.bias)

    def forward(self, x):
        b, t, _ = x.shape
        q, k, v = [m(x).view(b, t, self.h, self.d).transpose(1, 2) for m in (self.query, self.key, self.value)]
        o = F.scaled_dot_product_attention(q, k, v)          # softmax(q k^T / sqrt(key_dim)) v
        return self.output(o.transpose(1, 2).reshape(b, t, self.h * self.d))


class TransformerBlock(torch.nn.Module):
    """``layers.py:72-95``.  As written there, BOTH residuals add the block's INPUT: x = in + attn(norm_1(in)), then
    out = in + mlp(norm_2(x)); ``layer_norm_1`` is a BatchNormalization over the embedding axis."""

    def __init__(self, num_heads=12, embed_dim=768, mlp_ratio=4):
        super().__init__()
        self.layer_norm_1 = BatchNorm(embed_dim)
        self.attention = MultiHeadAttention(embed_dim, num_heads, embed_dim // num_heads)
        self.layer_norm_2 = torch.nn.LayerNorm(embed_dim, eps=_LN_EPS)
        self.mlp = torch.nn.Sequential(torch.nn.Linear(embed_dim, embed_dim * mlp_ratio), torch.nn.GELU(),
                                       torch.nn.Linear(embed_dim * mlp_ratio, embed_dim))

    def forward(self, inputs):
        x = self.layer_norm_1(inputs.transpose(1, 2)).transpose(1, 2)
        x = inputs + self.attention(x)
        return inputs + self.mlp(self.layer_norm_2(x))


class VisionTransformer(torch.nn.Module):
    """``layers.py:98-152`` (ViT-B/16 on 224 x 224, features taken after blocks 3, 6, 9, 12; classification skipped)"""

    def __init__(self, img_size=(224, 224), patch_size=16, embed_dim=768, mlp_ratio=4, hooks=(3, 6, 9, 12), num_heads=12):
        super().__init__()
        self.image_size = tuple(img_size)
        self.patch_embed = Conv2D(3, embed_dim, patch_size, strides=patch_size)
        self.grid_size = (img_size[0] // patch_size, img_size[1] // patch_size)
        n = self.grid_size[0] * self.grid_size[1] + 1
        self.cls_token = torch.nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embedding = torch.nn.Parameter(torch.randn(1, n, embed_dim) * 0.02)
        per = [hooks[0]] + [hooks[i] - hooks[i - 1] for i in range(1, len(hooks))]
        self.transformer_blocks = torch.nn.ModuleList(
            torch.nn.Sequential(*[TransformerBlock(num_heads, embed_dim, mlp_ratio) for _ in range(h)]) for h in per)

    def forward(self, images_nchw):
        x = self.patch_embed(images_nchw).flatten(2).transpose(1, 2)             # 'b h w c -> b (h w) c'
        x = torch.cat([self.cls_token.expand(x.shape[0], -1, -1), x], dim=1) + self.pos_embedding
        features = []
        for blocks in self.transformer_blocks:
            x = blocks(x)
            features.append(x)
        return x, features


class VisionTransformerEncoder(torch.nn.Module):
    """``layers.py:155-229``: DPT-style read-out of the four hooked token maps -> [B, n_features / 2, 112, 112] (NCHW)"""

    def __init__(self, img_size=(224, 224), patch_size=16, embed_dim=768, n_features=256, mlp_ratio=4,
                 hooks=(3, 6, 9, 12), features=(48, 96, 192, 384)):
        super().__init__()
        self.vit = VisionTransformer(img_size, patch_size, embed_dim, mlp_ratio, hooks)
        f = features
        self.post_process_1 = torch.nn.Sequential(Conv2D(embed_dim, f[0], 1), _ConvT(f[0], f[0], 4))
        self.post_process_2 = torch.nn.Sequential(Conv2D(embed_dim, f[1], 1), _ConvT(f[1], f[1], 2))
        self.post_process_3 = Conv2D(embed_dim, f[2], 1)
        self.post_process_4 = torch.nn.Sequential(Conv2D(embed_dim, f[3], 1), Conv2D(f[3], f[3], 3, strides=2, padding='same'))
        for i in range(4):
            setattr(self, f'conv_decode_{i + 1}', Conv2D(f[i], n_features, 3, padding='same', use_bias=False))
        self.output_conv = torch.nn.Sequential(torch.nn.ReLU(), Conv2D(4 * n_features, n_features, 3, padding='same'),
                                               torch.nn.ReLU(), Conv2D(n_features, n_features // 2, 3, padding='same'))

    def forward(self, images_nchw):
        _, feats = self.vit(images_nchw)
        gh, gw = self.vit.grid_size
        maps = [f[:, 1:].transpose(1, 2).reshape(f.shape[0], -1, gh, gw) for f in feats]      # drop the class token
        size = (gh * 8, gw * 8)                                  # 56 x2 = 28 x4 = 14 x8 = 7 x16 = 112: 8 x the token grid
        latents = []
        for i, m in enumerate(maps):
            y = getattr(self, f'conv_decode_{i + 1}')(getattr(self, f'post_process_{i + 1}')(m))
            latents.append(resize_bilinear(y, size))             # UpSampling2D(2 / 4 / 8 / 16, 'bilinear')
        return self.output_conv(torch.cat(latents, dim=1))


class _ConvT(torch.nn.ConvTranspose2d):
    """Keras ``Conv2DTranspose(filters, kernel_size=k, strides=k)`` (valid padding: non-overlapping k x k patches)"""

    def __init__(self, in_ch, filters, k):
        super().__init__(in_ch, filters, k, stride=k)
        torch.nn.init.xavier_uniform_(self.weight)
        torch.nn.init.zeros_(self.bias)


class VisualFeatures(torch.nn.Module):
    """``layers.py:232-259``: images [(b n), H, W, 3] in [0, 1] -> [(b n), H/2, W/2, n_features]: the ViT read-out (image
    resized to 224 x 224, result resized to half the original size) beside the stride-2 convolutional encoder"""

    def __init__(self, n_features=256, original_image_size=(480, 640)):
        super().__init__()
        self.conv_features = ConvolutionalEncoder(n_features)
        self.vision_transformer = VisionTransformerEncoder(n_features=n_features)
        self.half = (original_image_size[0] // 2, original_image_size[1] // 2)

    def forward(self, images_nhwc):
        x = _nchw(images_nhwc)
        latents = resize_bilinear(self.vision_transformer(resize_bilinear(x, (224, 224))), self.half)
        return torch.cat([_nhwc(latents), self.conv_features(images_nhwc)], dim=-1)


# ------------------------------------------------------------------------------------------------ CLIP fusion pyramid
class DoubleConv(torch.nn.Module):
    def __init__(self, in_ch, filters, activation='relu'):
        super().__init__()
        if activation not in ('relu', 'elu'):
            raise ValueError(f'activation {activation} not supported')
        self.conv_1 = Conv2D(in_ch, filters, 3, padding='same', use_bias=False, activation=activation)
        self.conv_2 = Conv2D(filters, filters, 3, padding='same', use_bias=False, activation=activation)

    def forward(self, x):
        return self.conv_2(self.conv_1(x))


class Up(torch.nn.Module):
    """``layers.py:437-456``: 2x bilinear upsample of x, CLIP level resized to ``shape``, concat, DoubleConv"""

    def __init__(self, shape, in_ch, clip_ch, filters, activation='relu'):
        super().__init__()
        self.shape = tuple(shape)
        self.double_conv = DoubleConv(in_ch + clip_ch, filters, activation)

    def forward(self, x, clip_x):
        x = resize_bilinear(x, (2 * x.shape[-2], 2 * x.shape[-1]))
        return self.double_conv(torch.cat([x, resize_bilinear(clip_x, self.shape)], dim=1))


class ConvFusion(torch.nn.Module):
    """``layers.py:459-477``: concat, activation, 1x1 conv (no bias)"""

    def __init__(self, in_ch, filters, activation='relu'):
        super().__init__()
        if activation not in ('relu', 'elu'):
            raise ValueError(f'activation {activation} not supported')
        self.activation = activation
        self.conv = Conv2D(in_ch, filters, 1, use_bias=False)

    def forward(self, x1, x2):
        return self.conv(_activate(torch.cat([x1, x2], dim=1), self.activation))


class MultiplyFusion(torch.nn.Module):
    """``layers.py:480-520``: x * tile(text) with text = Dense(filters, no bias)(clip_textuals) or its first ``filters``
    entries (``use_dense=False``: ``Slice``, whose tf.slice end index -1 on the batch axis means "all rows")"""

    def __init__(self, filters, use_dense=True):
        super().__init__()
        self.filters = filters
        self.dense = torch.nn.Linear(1024, filters, bias=False) if use_dense else None
        if self.dense is not None:
            torch.nn.init.xavier_uniform_(self.dense.weight)

    def forward(self, x, clip_textuals):
        t = self.dense(clip_textuals) if self.dense is not None else clip_textuals[:, :self.filters]
        return x * t[:, :, None, None]                     # the H x W tile of the reference is a broadcast


class CombineCLIPVisualV4(torch.nn.Module):
    """``layers.py:593-660``: U-Net-like fusion of the CLIP-RN50 pyramid, the text embedding and ``VisualFeatures`` ->
    [(b n), 480, 640, 256].  ``final_upsample=False`` stops before the last ``UpSampling2D`` (``:659``): the 240 x 320
    map the gather's low-resolution mode takes (same function of the same data; 4x smaller map and map gradient)."""
    _up3_filters = 128

    def __init__(self, use_dense=False, activation='relu', final_upsample=True):
        super().__init__()
        self.final_upsample = final_upsample
        self.conv = Conv2D(2048, 1024, 3, padding='same', use_bias=False, activation=activation)
        self.multiply_fusion_1 = MultiplyFusion(1024, use_dense)
        self.up_1 = Up((60, 80), 1024, 1024, 512, activation)
        self.multiply_fusion_2 = MultiplyFusion(512, use_dense)
        self.conv_fusion_1 = ConvFusion(512 + 256, 512, activation)
        self.up_2 = Up((120, 160), 512, 512, 256, activation)
        self.multiply_fusion_3 = MultiplyFusion(256, use_dense)
        self.conv_fusion_2 = ConvFusion(256 + 256, 256, activation)
        self.up_3 = Up((240, 320), 256, 256, self._up3_filters, activation)
        self.conv_fusion_3 = ConvFusion(self._up3_filters + 256, 256, activation)

    def forward(self, inputs):
        clip_outputs, visual_features, clip_textuals = inputs
        vis = _nchw(visual_features)                                       # [(BN) 256 240 320]
        vis_1, vis_2 = resize_bilinear(vis, (120, 160)), resize_bilinear(vis, (60, 80))
        l1, l2, l3, l4 = [_nchw(t) for t in clip_outputs[1:5]]
        x = self.conv(resize_bilinear(l4, (30, 40)))                       # [(BN) 1024 30 40]
        x = self.multiply_fusion_1(x, clip_textuals)
        x = self.up_1(x, l3)                                               # [(BN) 512 60 80]
        x = self.multiply_fusion_2(x, clip_textuals)
        x = self.conv_fusion_1(x, vis_2)
        x = self.up_2(x, l2)                                               # [(BN) 256 120 160]
        x = self.multiply_fusion_3(x, clip_textuals)
        x = self.conv_fusion_2(x, vis_1)
        x = self.up_3(x, l1)                                               # [(BN) 128|256 240 320]
        x = self.conv_fusion_3(x, vis)                                     # [(BN) 256 240 320]
        if self.final_upsample:
            x = resize_bilinear(x, (2 * x.shape[-2], 2 * x.shape[-1]))     # [(BN) 256 480 640]
        return _nhwc(x).contiguous()


class CombineCLIPVisualV3(CombineCLIPVisualV4):
    """``layers.py:523-590``: V4 with 256 instead of 128 channels out of the last ``Up``"""
    _up3_filters = 256


# ------------------------------------------------------------------------------------------------ checkpoint interchange
def keras_variables(module: torch.nn.Module, prefix: str = '') -> Dict[str, torch.Tensor]:
    """every variable of ``module`` in KERAS layout under its attribute path: Conv2D ``kernel`` HWIO, Conv2DTranspose
    ``kernel`` HWOI, Dense ``kernel`` [in, out], attention ``kernel`` [embed, heads, dim] (output: [heads, dim, embed]),
    BatchNormalization ``gamma / beta / moving_mean / moving_variance``, LayerNormalization ``gamma / beta``."""
    out = {}
    for name, m in module.named_modules():
        p = f'{prefix}{name}.' if name else prefix
        if isinstance(m, torch.nn.ConvTranspose2d):
            out[p + 'kernel'] = m.weight.permute(2, 3, 1, 0)
            out[p + 'bias'] = m.bias
        elif isinstance(m, torch.nn.Conv2d):
            out[p + 'kernel'] = m.weight.permute(2, 3, 1, 0)
            if m.bias is not None:
                out[p + 'bias'] = m.bias
        elif isinstance(m, MultiHeadAttention):
            for proj in ('query', 'key', 'value'):
                lin = getattr(m, proj)
                out[f'{p}{proj}.kernel'] = lin.weight.t().reshape(-1, m.h, m.d)
                out[f'{p}{proj}.bias'] = lin.bias.view(m.h, m.d)
            out[p + 'output.kernel'] = m.output.weight.t().reshape(m.h, m.d, -1)
            out[p + 'output.bias'] = m.output.bias
        elif isinstance(m, torch.nn.Linear):
            if any(name.endswith('.' + s) or name == s for s in ('query', 'key', 'value', 'output')) and \
                    isinstance(_parent(module, name), MultiHeadAttention):
                continue
            out[p + 'kernel'] = m.weight.t()
            if m.bias is not None:
                out[p + 'bias'] = m.bias
        elif isinstance(m, BatchNorm):
            out[p + 'gamma'], out[p + 'beta'] = m.weight, m.bias
            out[p + 'moving_mean'], out[p + 'moving_variance'] = m.running_mean, m.running_var
        elif isinstance(m, torch.nn.LayerNorm):
            out[p + 'gamma'], out[p + 'beta'] = m.weight, m.bias
        elif isinstance(m, VisionTransformer):
            out[p + 'cls_token'], out[p + 'pos_embedding'] = m.cls_token, m.pos_embedding
    return out


def _parent(root, name):
    mod = root
    for part in name.split('.')[:-1]:
        mod = getattr(mod, part)
    return mod


def load_keras_variables(module: torch.nn.Module, mapping, prefix: str = '', strict: bool = True) -> Sequence[str]:
    """copy ``mapping[name]`` (numpy arrays / tensors in Keras layout, e.g. an ``.npz`` written by
    ``tools/convert_tf_checkpoint.py``) into the module; returns the names that were missing (``strict``: raises)."""
    missing = []
    with torch.no_grad():
        for name, view in keras_variables(module, prefix).items():
            if name not in mapping:
                missing.append(name)
                continue
            src = torch.as_tensor(mapping[name], dtype=view.dtype)
            if tuple(src.shape) != tuple(view.shape):
                raise ValueError(f'{name}: checkpoint shape {tuple(src.shape)} != {tuple(view.shape)}')
            view.copy_(src.to(view.device))           # views into the torch parameters: written in place
    if missing and strict:
        raise KeyError(f'variables missing from the checkpoint: {missing[:5]}{"..." if len(missing) > 5 else ""}')
    return missing
