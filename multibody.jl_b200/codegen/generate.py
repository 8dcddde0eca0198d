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


if __name__ == "__main__":
    main(sys.argv[1:])
