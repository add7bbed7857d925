"""Drop-in ``AVENet`` / ``FullModel`` / ``HardWayAttention`` with the reference's call
signatures (``/root/reference/model.py``), the head running in libavtubes.

    AVENet(args, pretrained).forward(image, audio) -> (A, logits, Pos, Neg)     model.py:112-154
    FullModel(args).forward(audio, video)          -> (A, logits)               model.py:26-36
    HardWayAttention().forward(audio_features, video_features) -> (A, logits)   model.py:46-60

Differences, all deliberate: no ``.cuda()`` calls inside forward (inputs decide the device);
no global ``activation`` dict (the racy forward hook of ``model.py:11-15`` becomes a direct
call of the 3-D trunk); ``return_weighted=True`` gives HEAD's 5-tuple
``(A, logits, weighted_A, Pos, Neg)`` (``model.py:154``).
"""
from __future__ import annotations

import torch
from torch import nn

from .backbones import resnet18_2d, resnet18_3d
from .head import HeadConfig, audio_normalize, hardway_head, hardway_loss


class AVENet(nn.Module):
    def __init__(self, args, pretrained=False, precision: str = "f16", return_weighted: bool = False):
        super().__init__()
        self.imgnet = resnet18_2d("vision")
        self.audnet = resnet18_2d("audio")
        self.avgpool = nn.AdaptiveMaxPool2d((1, 1))          # sic: a max-pool, model.py:96
        self.epsilon = args.epsilon
        self.epsilon2 = args.epsilon2
        self.tau = 0.03
        self.trimap = args.tri_map
        self.Neg = args.Neg
        self.precision = precision
        self.return_weighted = return_weighted
        for m in self.modules():                             # model.py:104-110
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, (nn.BatchNorm2d, nn.GroupNorm)):
                nn.init.normal_(m.weight, mean=1, std=0.02)
                nn.init.constant_(m.bias, 0)

    @property
    def head_config(self) -> HeadConfig:
        return HeadConfig(eps=self.epsilon, eps2=self.epsilon2, tau=self.tau, temp=0.07,
                          mask_scale=100.0, trimap=bool(self.trimap), use_neg=bool(self.Neg),
                          precision=self.precision)

    def features(self, image, audio):
        """Backbones + the audio max-pool: (img [B,512,14,14], aud [B,512]), un-normalised."""
        B = image.shape[0]
        img = self.imgnet(image)
        aud = self.avgpool(self.audnet(audio)).view(B, -1)
        return img, aud

    def forward(self, image, audio):
        img, aud = self.features(image, audio)
        B, _, H, W = img.shape
        A, logits, Pos, Neg, wA, _ = hardway_head(img, aud, self.head_config,
                                                  return_weighted=self.return_weighted)
        A, Pos, Neg = (t.view(B, 1, H, W) for t in (A, Pos, Neg))
        if self.return_weighted:
            return A, logits, wA.view(B, H, W), Pos, Neg
        return A, logits, Pos, Neg

    def loss(self, image, audio):
        """``CrossEntropyLoss()(logits, zeros)`` of the forward, fused (train_hardway_1frame.py:129-131)."""
        img, aud = self.features(image, audio)
        return hardway_loss(img, aud, self.head_config)


class HardWayAttention(nn.Module):
    """model.py:38-60.  ``audio_features [N_a,512]`` (any scale: the head re-normalises, which
    is the identity on the reference's already-normalised rows), ``video_features
    [b,512,T,14,14]``.  ``N_a == 1`` broadcasts against the b*T frames exactly like the
    reference's einsum does at test time (train_3D.py:153-154).  ``N_a == b`` with T > 1 is the
    de-duplicated tube mode: pass the b distinct clip embeddings instead of
    repeating each T times (train_3D.py:129-130) and get the same ``(A, logits)`` -- logits
    ``[(bT), (bT)+2]`` -- at 1/T of the similarity flops (and run the audio backbone on b
    spectrograms instead of bT)."""

    def __init__(self, precision: str = "f16"):
        super().__init__()
        self.epsilon, self.epsilon2, self.tau = 0.65, 0.4, 0.03
        self.precision = precision

    def forward(self, audio_features, video_features):
        b, c, t, h, w = video_features.shape
        n = b * t
        v = video_features.permute(0, 2, 1, 3, 4).reshape(n, c, h * w)      # 'b c t h w -> (b t) c (h w)'
        aud = audio_features
        if aud.shape[0] == 1 and n > 1:
            aud = aud.expand(n, -1)
        cfg = HeadConfig(eps=self.epsilon, eps2=self.epsilon2, tau=self.tau, precision=self.precision)
        col_mult = t if (aud.shape[0] == b and t > 1 and b > 1) else 1
        A, logits, _, _, _, _ = hardway_head(v.contiguous(), aud.contiguous(), cfg, col_mult=col_mult)
        return A.view(n, 1, h, w), logits


class FullModel(nn.Module):
    def __init__(self, args=None, precision: str = "f16"):
        super().__init__()
        self.vidnet = resnet18_3d(n_classes=1039)
        self.audnet = resnet18_2d("audio")
        self.avgpool = nn.AdaptiveMaxPool2d((1, 1))
        self.attention = HardWayAttention(precision)

    def forward(self, audio, video):
        B = audio.shape[0]
        aud = self.avgpool(self.audnet(audio)).view(B, -1)
        vid = self.vidnet(video).detach()                    # model.py:14: the hook detaches
        return self.attention(aud, vid)
