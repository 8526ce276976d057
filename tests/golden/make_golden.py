"""Generates tests/golden/*.npz by running the UNMODIFIED reference CPU path (oracle/_ref/libyolo_ref.so built
from /root/reference by oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py

Each fixture holds: the cfg text, the synthetic weights seed parameters, the input image, selected per-layer
outputs of the reference's NetworkPredict (fused BN, batch 1), and its GetNetworkBoxes / NmsSort results.
The fixtures are small (micro net 64x64) so they can live in git; they pin the oracle and the CUDA path on
machines where /root/reference does not exist."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from darknet_b200 import cfg as cfgmod  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.netutil import materialise  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(tag, name, kw, head_bias, head_gain, thresh):
    text, cpath, wpath = materialise(name, head_bias=head_bias, head_gain=head_gain, **kw)
    x = cfgmod.synth_images(1, 3, 64, 64)
    ref = pyoracle.RefNet(cpath, wpath, train=False, fuse=True)
    ref.predict(x)
    d = {"cfg": np.array(text), "name": np.array(name), "kw": np.array(repr(kw)), "head_bias": np.array(head_bias),
         "head_gain": np.array(head_gain), "thresh": np.array(thresh), "x": x.astype(np.float16)}
    # the image is stored as fp16 to keep the fixture small; the test regenerates it from the seed and checks
    for i in range(ref.n):
        d[f"out{i}"] = ref.layer_output(i).astype(np.float32)
    pre7, prep = ref.boxes(thresh, 0.0)
    post7, postp = ref.boxes(thresh, 0.45)
    d.update(pre7=pre7, prep=prep, post7=post7, postp=postp)
    ref.close()
    np.savez_compressed(os.path.join(HERE, tag + ".npz"), **d)
    print(tag, "layers", ref.n, "dets", pre7.shape[0])


if __name__ == "__main__":
    make("micro_leaky", "micro", {}, -3.0, 40.0, 0.25)
    make("micro_mish", "micro", {"act": "mish"}, -3.0, 40.0, 0.25)
    make("micro_gaussian", "micro", {"gaussian": True}, -3.0, 40.0, 0.25)
