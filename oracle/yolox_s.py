"""ORACLE (test infrastructure only) -- CPU restatement of YOLOX-S.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product path never does.

The reference imports the network from an un-vendored checkout of
Megvii-BaseDetection/YOLOX (no version pin; weights release 0.1.1rc0 is named at
/root/reference/main.py:386; call sites /root/reference/src/yolox_detector.py:36-37,
131-137,306-314,378-386).  The published architecture (yolox/models/{darknet,
network_blocks,yolo_pafpn,yolo_head,yolox}.py, yolox/exp/default/yolox_s.py) is
restated here in plain fp32 PyTorch; module / parameter names follow upstream so a
real ``yolox_s.pth`` state dict loads with strict=True.

PARITY UNPINNED: the reference ships no golden vectors for this path
and upstream YOLOX is absent from this image; structure is cross-checked by
parameter / FLOP count only (8.97 M params, 26.69 GFLOP at 640x640).
"""
import math
import torch
import torch.nn as nn


class BaseConv(nn.Module):
    """Conv2d(bias=False) -> BN(eps=1e-3) -> SiLU."""

    def __init__(self, cin, cout, k, s):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, s, (k - 1) // 2, bias=False)
        self.bn = nn.BatchNorm2d(cout, eps=1e-3, momentum=0.03)
        self.act = nn.SiLU()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class Bottleneck(nn.Module):
    def __init__(self, cin, cout, shortcut=True, expansion=1.0):
        super().__init__()
        h = int(cout * expansion)
        self.conv1 = BaseConv(cin, h, 1, 1)
        self.conv2 = BaseConv(h, cout, 3, 1)
        self.use_add = shortcut and cin == cout

    def forward(self, x):
        y = self.conv2(self.conv1(x))
        return y + x if self.use_add else y


class CSPLayer(nn.Module):
    def __init__(self, cin, cout, n=1, shortcut=True):
        super().__init__()
        h = cout // 2
        self.conv1 = BaseConv(cin, h, 1, 1)
        self.conv2 = BaseConv(cin, h, 1, 1)
        self.conv3 = BaseConv(2 * h, cout, 1, 1)
        self.m = nn.Sequential(*[Bottleneck(h, h, shortcut, 1.0) for _ in range(n)])

    def forward(self, x):
        x1 = self.m(self.conv1(x))
        x2 = self.conv2(x)
        return self.conv3(torch.cat((x1, x2), dim=1))


class SPPBottleneck(nn.Module):
    def __init__(self, cin, cout, ks=(5, 9, 13)):
        super().__init__()
        h = cin // 2
        self.conv1 = BaseConv(cin, h, 1, 1)
        self.m = nn.ModuleList([nn.MaxPool2d(k, 1, k // 2) for k in ks])
        self.conv2 = BaseConv(h * (len(ks) + 1), cout, 1, 1)

    def forward(self, x):
        x = self.conv1(x)
        return self.conv2(torch.cat([x] + [m(x) for m in self.m], dim=1))


class Focus(nn.Module):
    def __init__(self, cin, cout, k=3):
        super().__init__()
        self.conv = BaseConv(cin * 4, cout, k, 1)

    def forward(self, x):
        tl = x[..., ::2, ::2]
        tr = x[..., ::2, 1::2]
        bl = x[..., 1::2, ::2]
        br = x[..., 1::2, 1::2]
        return self.conv(torch.cat((tl, bl, tr, br), dim=1))


class CSPDarknet(nn.Module):
    def __init__(self, dep=0.33, wid=0.5):
        super().__init__()
        bc = int(wid * 64)
        bd = max(round(dep * 3), 1)
        self.stem = Focus(3, bc, 3)
        self.dark2 = nn.Sequential(BaseConv(bc, bc * 2, 3, 2), CSPLayer(bc * 2, bc * 2, bd))
        self.dark3 = nn.Sequential(BaseConv(bc * 2, bc * 4, 3, 2), CSPLayer(bc * 4, bc * 4, bd * 3))
        self.dark4 = nn.Sequential(BaseConv(bc * 4, bc * 8, 3, 2), CSPLayer(bc * 8, bc * 8, bd * 3))
        self.dark5 = nn.Sequential(
            BaseConv(bc * 8, bc * 16, 3, 2),
            SPPBottleneck(bc * 16, bc * 16),
            CSPLayer(bc * 16, bc * 16, bd, shortcut=False),
        )

    def forward(self, x):
        x = self.stem(x)
        x = self.dark2(x)
        d3 = self.dark3(x)
        d4 = self.dark4(d3)
        d5 = self.dark5(d4)
        return d3, d4, d5


class YOLOPAFPN(nn.Module):
    def __init__(self, dep=0.33, wid=0.5, in_channels=(256, 512, 1024)):
        super().__init__()
        self.backbone = CSPDarknet(dep, wid)
        c0, c1, c2 = [int(c * wid) for c in in_channels]
        n = round(3 * dep)
        self.upsample = nn.Upsample(scale_factor=2, mode="nearest")
        self.lateral_conv0 = BaseConv(c2, c1, 1, 1)
        self.C3_p4 = CSPLayer(2 * c1, c1, n, False)
        self.reduce_conv1 = BaseConv(c1, c0, 1, 1)
        self.C3_p3 = CSPLayer(2 * c0, c0, n, False)
        self.bu_conv2 = BaseConv(c0, c0, 3, 2)
        self.C3_n3 = CSPLayer(2 * c0, c1, n, False)
        self.bu_conv1 = BaseConv(c1, c1, 3, 2)
        self.C3_n4 = CSPLayer(2 * c1, c2, n, False)

    def forward(self, x):
        x2, x1, x0 = self.backbone(x)
        fpn_out0 = self.lateral_conv0(x0)
        f_out0 = self.C3_p4(torch.cat([self.upsample(fpn_out0), x1], 1))
        fpn_out1 = self.reduce_conv1(f_out0)
        pan_out2 = self.C3_p3(torch.cat([self.upsample(fpn_out1), x2], 1))
        pan_out1 = self.C3_n3(torch.cat([self.bu_conv2(pan_out2), fpn_out1], 1))
        pan_out0 = self.C3_n4(torch.cat([self.bu_conv1(pan_out1), fpn_out0], 1))
        return pan_out2, pan_out1, pan_out0


class YOLOXHead(nn.Module):
    def __init__(self, num_classes=80, wid=0.5, strides=(8, 16, 32), in_channels=(256, 512, 1024)):
        super().__init__()
        self.num_classes = num_classes
        self.strides = strides
        h = int(256 * wid)
        self.stems = nn.ModuleList([BaseConv(int(c * wid), h, 1, 1) for c in in_channels])
        self.cls_convs = nn.ModuleList(
            [nn.Sequential(BaseConv(h, h, 3, 1), BaseConv(h, h, 3, 1)) for _ in in_channels])
        self.reg_convs = nn.ModuleList(
            [nn.Sequential(BaseConv(h, h, 3, 1), BaseConv(h, h, 3, 1)) for _ in in_channels])
        self.cls_preds = nn.ModuleList([nn.Conv2d(h, num_classes, 1) for _ in in_channels])
        self.reg_preds = nn.ModuleList([nn.Conv2d(h, 4, 1) for _ in in_channels])
        self.obj_preds = nn.ModuleList([nn.Conv2d(h, 1, 1) for _ in in_channels])

    def initialize_biases(self, prior_prob):
        b = -math.log((1 - prior_prob) / prior_prob)
        for conv in list(self.cls_preds) + list(self.obj_preds):
            nn.init.constant_(conv.bias, b)

    def forward(self, feats):
        outs, grids, strides = [], [], []
        for k, x in enumerate(feats):
            x = self.stems[k](x)
            cls_feat = self.cls_convs[k](x)
            reg_feat = self.reg_convs[k](x)
            o = torch.cat([self.reg_preds[k](reg_feat),
                           self.obj_preds[k](reg_feat).sigmoid(),
                           self.cls_preds[k](cls_feat).sigmoid()], 1)
            hh, ww = o.shape[-2:]
            yv, xv = torch.meshgrid(torch.arange(hh), torch.arange(ww), indexing="ij")
            grids.append(torch.stack((xv, yv), 2).view(1, -1, 2).to(o.dtype))
            strides.append(torch.full((1, hh * ww, 1), float(self.strides[k]), dtype=o.dtype))
            outs.append(o.flatten(2))
        out = torch.cat(outs, 2).permute(0, 2, 1)
        g = torch.cat(grids, 1)
        s = torch.cat(strides, 1)
        return torch.cat([(out[..., 0:2] + g) * s, torch.exp(out[..., 2:4]) * s, out[..., 4:]], -1)


class YOLOX(nn.Module):
    def __init__(self, dep=0.33, wid=0.5, num_classes=80):
        super().__init__()
        self.backbone = YOLOPAFPN(dep, wid)
        self.head = YOLOXHead(num_classes, wid)

    def forward(self, x):
        return self.head(self.backbone(x))


class Exp:
    """Stand-in for yolox.exp.get_exp(None, 'yolox-s') (only what the reference reads)."""

    def __init__(self, name="yolox-s"):
        table = {"yolox-s": (0.33, 0.50), "yolox-m": (0.67, 0.75), "yolox-l": (1.0, 1.0),
                 "yolox-tiny": (0.33, 0.375), "yolox-nano": (0.33, 0.25)}
        self.depth, self.width = table[name]
        self.num_classes = 80
        self.test_size = (640, 640)
        self.test_conf = 0.01
        self.nmsthre = 0.65

    def get_model(self):
        m = YOLOX(self.depth, self.width, self.num_classes)
        m.head.initialize_biases(1e-2)
        return m


def get_exp(exp_file=None, exp_name="yolox-s"):
    return Exp(exp_name)
