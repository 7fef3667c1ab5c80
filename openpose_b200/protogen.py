"""Deploy-prototxt generators for the four OpenPose networks (BODY_25, COCO, hand, face).

The reference ships these graphs as prototxt files (models/pose/body_25/pose_deploy.prototxt,
models/pose/coco/pose_deploy_linevec.prototxt, models/hand/pose_deploy.prototxt,
models/face/pose_deploy.prototxt) and downloads the trained weights. Neither travels to the GPU box, so the
harness regenerates the graphs from their architecture description. The generated text is
a normal Caffe deploy prototxt: NetB200 loads it exactly like a user-supplied file. tests/test_protogen.py
checks layer-by-layer (name, type, bottoms, tops, parameters) that the generated graphs equal the reference's
files whenever /root/reference is present, and against committed signatures otherwise.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

MODELS = ("body_25", "coco", "hand", "face")


class _Builder:
    def __init__(self, name: Optional[str], input_dim=(1, 3, 16, 16)):
        self.out: List[str] = []
        if name:
            self.out.append('name: "%s"' % name)
        self.out.append('input: "image"')
        for d in input_dim:
            self.out.append("input_dim: %d" % d)

    def conv(self, name: str, bottom: str, cout: int, k: int, top: Optional[str] = None) -> str:
        top = top or name
        self.out.append('layer {\n  name: "%s"\n  type: "Convolution"\n  bottom: "%s"\n  top: "%s"\n'
                        "  convolution_param {\n    num_output: %d\n    pad: %d\n    kernel_size: %d\n  }\n}"
                        % (name, bottom, top, cout, k // 2, k))
        return top

    def act(self, kind: str, name: str, blob: str) -> None:
        self.out.append('layer {\n  name: "%s"\n  type: "%s"\n  bottom: "%s"\n  top: "%s"\n}' % (name, kind, blob, blob))

    def pool(self, name: str, bottom: str) -> str:
        self.out.append('layer {\n  name: "%s"\n  type: "Pooling"\n  bottom: "%s"\n  top: "%s"\n'
                        "  pooling_param {\n    pool: MAX\n    kernel_size: 2\n    stride: 2\n  }\n}" % (name, bottom, name))
        return name

    def concat(self, name: str, bottoms: List[str], top: Optional[str] = None) -> str:
        top = top or name
        b = "".join('  bottom: "%s"\n' % x for x in bottoms)
        self.out.append('layer {\n  name: "%s"\n  type: "Concat"\n%s  top: "%s"\n  concat_param {\n    axis: 1\n  }\n}'
                        % (name, b, top))
        return top

    def text(self) -> str:
        return "\n".join(self.out) + "\n"


def _vgg_trunk(b: _Builder, relu_name, pool_name, n4: int) -> str:
    """conv1_1 .. conv4_{n4}: VGG-19 front, 3x3 convs + ReLU, 2x2 max-pools after blocks 1-3."""
    x = "image"
    for block, (width, count) in enumerate([(64, 2), (128, 2), (256, 4), (512, n4)], start=1):
        for i in range(1, count + 1):
            name = "conv%d_%d" % (block, i)
            x = b.conv(name, x, width, 3)
            b.act("ReLU", relu_name(name, block, i), x)
        if block <= 3:
            x = b.pool(pool_name(block), x)
    return x


def body_25() -> str:
    b = _Builder("OpenPose - BODY_25")
    x = "image"
    for block, (width, count) in enumerate([(64, 2), (128, 2), (256, 4)], start=1):
        for i in range(1, count + 1):
            x = b.conv("conv%d_%d" % (block, i), x, width, 3)
            b.act("ReLU", "relu%d_%d" % (block, i), x)
        x = b.pool("pool%d_stage1" % block, x)
    x = b.conv("conv4_1", x, 512, 3)
    b.act("ReLU", "relu4_1", x)
    x = b.conv("conv4_2", x, 512, 3)
    b.act("PReLU", "prelu4_2", x)
    x = b.conv("conv4_3_CPM", x, 256, 3)
    b.act("PReLU", "prelu4_3_CPM", x)
    feat = b.conv("conv4_4_CPM", x, 128, 3)
    b.act("PReLU", "prelu4_4_CPM", feat)

    def stage(s: int, branch: str, src: str, width: int, mid: int, cout: int) -> str:
        x = src
        for blk in range(1, 6):
            outs = []
            for j in range(3):
                name = "Mconv%d_stage%d_%s_%d" % (blk, s, branch, j)
                x = b.conv(name, x, width, 3)
                b.act("PReLU", "Mprelu%d_stage%d_%s_%d" % (blk, s, branch, j), x)
                outs.append(x)
            x = b.concat("Mconv%d_stage%d_%s_concat" % (blk, s, branch), outs)
        x = b.conv("Mconv6_stage%d_%s" % (s, branch), x, mid, 1)
        b.act("PReLU", "Mprelu6_stage%d_%s" % (s, branch), x)
        return b.conv("Mconv7_stage%d_%s" % (s, branch), x, cout, 1)

    paf = stage(0, "L2", feat, 96, 256, 52)
    for s in (1, 2, 3):
        src = b.concat("concat_stage%d_L2" % s, [feat, paf])
        paf = stage(s, "L2", src, 128, 512, 52)
    src = b.concat("concat_stage0_L1", [feat, paf])
    hm = stage(0, "L1", src, 96, 256, 26)
    src = b.concat("concat_stage1_L1", [feat, hm, paf])
    hm = stage(1, "L1", src, 128, 512, 26)
    b.concat("net_output", [hm, paf])
    return b.text()


def coco() -> str:
    b = _Builder(None, (1, 3, 1, 1))
    x = _vgg_trunk(b, lambda n, blk, i: "relu%d_%d" % (blk, i), lambda blk: "pool%d_stage1" % blk, 2)
    x = b.conv("conv4_3_CPM", x, 256, 3)
    b.act("ReLU", "relu4_3_CPM", x)
    feat = b.conv("conv4_4_CPM", x, 128, 3)
    b.act("ReLU", "relu4_4_CPM", feat)
    # stage 1: the two branches are interleaved layer by layer in the file
    l1 = l2 = feat
    for i, (cout, k) in enumerate([(128, 3), (128, 3), (128, 3), (512, 1)], start=1):
        l1 = b.conv("conv5_%d_CPM_L1" % i, l1, cout, k)
        b.act("ReLU", "relu5_%d_CPM_L1" % i, l1)
        l2 = b.conv("conv5_%d_CPM_L2" % i, l2, cout, k)
        b.act("ReLU", "relu5_%d_CPM_L2" % i, l2)
    l1 = b.conv("conv5_5_CPM_L1", l1, 38, 1)
    l2 = b.conv("conv5_5_CPM_L2", l2, 19, 1)
    for s in range(2, 7):
        src = b.concat("concat_stage%d" % s, [l1, l2, feat])
        l1 = l2 = src
        for i in range(1, 7):
            cout, k = (128, 7) if i <= 5 else (128, 1)
            l1 = b.conv("Mconv%d_stage%d_L1" % (i, s), l1, cout, k)
            b.act("ReLU", "Mrelu%d_stage%d_L1" % (i, s), l1)
            l2 = b.conv("Mconv%d_stage%d_L2" % (i, s), l2, cout, k)
            b.act("ReLU", "Mrelu%d_stage%d_L2" % (i, s), l2)
        l1 = b.conv("Mconv7_stage%d_L1" % s, l1, 38, 1)
        l2 = b.conv("Mconv7_stage%d_L2" % s, l2, 19, 1)
    b.concat("concat_stage7", [l2, l1], top="net_output")
    return b.text()


def _cpm_single_branch(b: _Builder, trunk_out: str, cout: int, relu_cpm, concat_name, relu_stage) -> None:
    feat = b.conv("conv5_3_CPM", trunk_out, 128, 3)
    b.act("ReLU", relu_cpm("conv5_3_CPM"), feat)
    x = b.conv("conv6_1_CPM", feat, 512, 1)
    b.act("ReLU", relu_cpm("conv6_1_CPM"), x)
    prev = b.conv("conv6_2_CPM", x, cout, 1)
    for s in range(2, 7):
        x = b.concat(concat_name(s), [prev, feat])
        for i in range(1, 7):
            c, k = (128, 7) if i <= 5 else (128, 1)
            x = b.conv("Mconv%d_stage%d" % (i, s), x, c, k)
            b.act("ReLU", relu_stage(i, s), x)
        prev = b.conv("Mconv7_stage%d" % s, x, cout, 1, top="net_output" if s == 6 else None)


def hand() -> str:
    b = _Builder(None, (1, 3, 1, 1))
    x = _vgg_trunk(b, lambda n, blk, i: "relu%d_%d" % (blk, i), lambda blk: "pool%d_stage1" % blk, 4)
    for i in (1, 2):
        x = b.conv("conv5_%d" % i, x, 512, 3)
        b.act("ReLU", "relu5_%d" % i, x)
    relu_cpm = {"conv5_3_CPM": "relu5_4_stage1_3", "conv6_1_CPM": "relu6_4_stage1_1"}
    _cpm_single_branch(b, x, 22, lambda n: relu_cpm[n], lambda s: "concat_stage%d" % s,
                       lambda i, s: "Mrelu1_%d_stage%d_%d" % (i + 1, s, i))
    return b.text()


def face() -> str:
    b = _Builder(None, (1, 3, 1, 1))
    x = _vgg_trunk(b, lambda n, blk, i: n + "_re", lambda blk: "pool%d" % blk, 4)
    for i in (1, 2):
        x = b.conv("conv5_%d" % i, x, 512, 3)
        b.act("ReLU", "conv5_%d_re" % i, x)
    _cpm_single_branch(b, x, 71, lambda n: n + "_re", lambda s: "features_in_stage_%d" % s,
                       lambda i, s: "Mconv%d_stage%d_re" % (i, s))
    return b.text()


def generate(model: str) -> str:
    return {"body_25": body_25, "coco": coco, "hand": hand, "face": face}[model]()


REFERENCE_FILES = {
    "body_25": "models/pose/body_25/pose_deploy.prototxt",
    "coco": "models/pose/coco/pose_deploy_linevec.prototxt",
    "hand": "models/hand/pose_deploy.prototxt",
    "face": "models/face/pose_deploy.prototxt",
}
OUTPUT_CHANNELS = {"body_25": 78, "coco": 57, "hand": 22, "face": 71}
CONV_GFLOP_368x656 = {"body_25": 287.26, "coco": 484.63}
