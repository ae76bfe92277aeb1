"""Which objects of one pix2pix step end up in reference cycles (diagnostic)."""
import collections, gc, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import bench_extra as B

def timed(fn, steps, warm=3):
    for i in range(3):
        fn()
    torch.cuda.synchronize()
    gc.collect()
    gc.set_debug(gc.DEBUG_SAVEALL)
    fn()
    torch.cuda.synchronize()
    n = gc.collect()
    cnt = collections.Counter(type(o).__name__ for o in gc.garbage)
    print("unreachable", n, cnt.most_common(25))
    big = [o for o in gc.garbage if isinstance(o, torch.Tensor) and o.numel() > 1 << 16]
    print("large tensors in garbage:", len(big), sum(t.numel() * t.element_size() for t in big) / 2**30, "GiB")
    for o in gc.garbage:
        if type(o).__name__ not in ("tuple", "PyCArrayType", "getset_descriptor", "StgDict", "Tensor", "dict", "list", "cell", "function"):
            refs = [type(r).__name__ for r in gc.get_referents(o)][:8]
            print("  ", type(o).__name__, str(o)[:80].replace("\n", " "), "->", refs)
    gc.set_debug(0); gc.garbage.clear()
    return 1.0
B.timed = timed
B.pix2pix(8, 1)
