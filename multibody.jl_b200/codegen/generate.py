"""Regenerate csrc/generated/mbm_<model>.h for every model in the zoo.

    python multibody.jl_b200/codegen/generate.py [model ...]
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import mbpkg  # noqa: E402

mbpkg.load()
from multibody_jl_b200.mechanics.zoo import MODELS, get_model  # noqa: E402
from multibody_jl_b200.codegen.model_emit import ModelCode, emit_model_header  # noqa: E402


def main(argv):
    names = argv or list(MODELS)
    outdir = os.path.join(os.path.dirname(HERE), "csrc", "generated")
    os.makedirs(outdir, exist_ok=True)
    for name in names:
        sys_ = get_model(name)
        pivots = sys_.meta.get("pivots")
        code = ModelCode(sys_, pivots=pivots)
        text = emit_model_header(sys_, code)
        path = os.path.join(outdir, f"mbm_{name}.h")
        old = open(path).read() if os.path.exists(path) else None
        if old != text:
            with open(path, "w") as f:
                f.write(text)
        print(f"{name}: n={sys_.n} m={sys_.m} nnz={code.nnz} nlu={code.nlu} F_rhs={code.F_rhs} "
              f"F_jac={code.F_jac} F_factor={code.F_factor} F_solve={code.F_solve} -> {path}")


def sources_hash() -> str:
    """Hash of everything the generated headers depend on: the model front end and the code generator."""
    import hashlib
    h = hashlib.sha256()
    pkg = os.path.dirname(HERE)
    for sub in ("mechanics", "codegen"):
        for fn in sorted(os.listdir(os.path.join(pkg, sub))):
            if fn.endswith(".py"):
                h.update(fn.encode())
                h.update(open(os.path.join(pkg, sub, fn), "rb").read())
    for k in ("MB_ROBOT_SLOTS", "MB_ROBOT_RTARGET", "MB_ROBOT_STAGE_FACTOR"):      # generator knobs
        h.update(f"{k}={os.environ.get(k, '')}".encode())
    return h.hexdigest()


def ensure_generated(verbose: bool = False) -> bool:
    """csrc/generated/ is build output (not committed): regenerate when a header is missing or the sources changed.
    Returns True when it regenerated."""
    outdir = os.path.join(os.path.dirname(HERE), "csrc", "generated")
    stamp = os.path.join(outdir, ".stamp")
    want = sources_hash()
    have = open(stamp).read().strip() if os.path.exists(stamp) else ""
    missing = [m for m in MODELS if not os.path.exists(os.path.join(outdir, f"mbm_{m}.h"))]
    if have == want and not missing:
        return False
    if verbose:
        print("regenerating csrc/generated (%s)" % ("missing: " + ", ".join(missing) if missing else "sources changed"))
    import contextlib, io
    with contextlib.redirect_stdout(sys.stdout if verbose else io.StringIO()):
        main([])
    with open(stamp, "w") as f:
        f.write(want + "\n")
    return True


if __name__ == "__main__":
    if sys.argv[1:] == ["--ensure"]:
        ensure_generated(verbose=True)
    else:
        main(sys.argv[1:])
