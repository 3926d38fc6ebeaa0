"""Source-level guard for programmatic dependent launch: a kernel launched through launch_chained() may become resident
before the previous kernel of its stream has finished, so it MUST order itself with grid_dep_wait() / grid_dep_sync()
before touching global memory (csrc/common/ptx.cuh, csrc/common/host.h).  A kernel that forgets is a silent data race; this
test reads the sources and fails if any chained launch names a kernel without the wait."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# launches through a function pointer: which kernel template the pointer holds
POINTERS = {"plan->fn": {"igemm_conv.cu": "igemm_conv_kernel", "igemm_wgrad.cu": "igemm_wgrad_kernel"},
            "kernel": {"loss_optim.cu": "head_sigmoid_bwd_kernel"}}


def _kernel_bodies(src: str):
    """name -> body text of every __global__ function in `src` (brace matching from the opening brace)."""
    out = {}
    src = re.sub(r"__launch_bounds__\([^)]*\)", "", src)
    for m in re.finditer(r"__global__[^;{(]*?\b([A-Za-z_]\w*)\s*\(", src):
        name = m.group(1)
        # closing parenthesis of the parameter list, then the body's opening brace
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(src[i], 0)
            i += 1
        start = src.index("{", i)
        j, depth = start + 1, 1
        while depth:
            depth += {"{": 1, "}": -1}.get(src[j], 0)
            j += 1
        out[name] = src[start:j]
    return out


def test_every_chained_launch_targets_a_kernel_that_waits_for_its_predecessor():
    files = sorted(glob.glob(os.path.join(ROOT, "csrc", "**", "*.cu"), recursive=True))
    bodies, launches = {}, []
    for f in files:
        src = open(f).read()
        bodies.update(_kernel_bodies(src))
        for m in re.finditer(r"launch_chained\(\s*([A-Za-z_][\w>\-]*)", src):
            name = m.group(1)
            if name in POINTERS:
                name = POINTERS[name][os.path.basename(f)]
            launches.append((os.path.basename(f), name))
    assert len(launches) >= 40, launches  # conv, wgrad, the reductions and every elementwise / loss / optimizer kernel
    missing = [(f, k) for f, k in launches if k not in bodies or not re.search(r"grid_dep_(wait|sync)\(\)", bodies[k])]
    assert not missing, f"chained launches of kernels without grid_dep_wait(): {missing}"
    # and nothing may read or write global memory before the wait in the two tcgen05 kernels: the wait sits right after
    # the set-up barrier, before the role dispatch
    for k in ("igemm_conv_kernel", "igemm_wgrad_kernel"):
        body = bodies[k]
        assert body.index("grid_dep_wait()") < body.index("if (warp == 0)"), k
