"""Builds the C part of the oracle (oracle/_build/liboracle.so).  Test infrastructure only.

There is no `oracle/_ref`: the reference is Rust over tch/libtorch plus the un-vendored tch-nms
crate; neither cargo nor rustc exists in this image, so no file of the reference's hot path can
be compiled here.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def build(force: bool = False) -> str:
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "liboracle.so")
    srcs = [os.path.join(HERE, "matcher.c")]
    if not force and os.path.exists(out) and all(
            os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs):
        return out
    cmd = ["gcc", "-O2", "-std=c11", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-o", out,
           *srcs, "-lm"]
    subprocess.run(cmd, check=True)
    return out


if __name__ == "__main__":
    print(build(force=True))
