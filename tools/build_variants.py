"""Build tuning variants of libgsr_b200.so into cf-3dgs_b200/variants/ (development aid)."""
import importlib.util, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("gsr_build", os.path.join(ROOT, "cf-3dgs_b200", "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
vdir = os.path.join(ROOT, "cf-3dgs_b200", "variants"); os.makedirs(vdir, exist_ok=True)
for spec_ in sys.argv[1:]:
    name, _, defs = spec_.partition(":")
    extra = ["-D" + d for d in defs.split(",") if d]
    print(name, extra, flush=True)
    b.build(force=True, extra=extra, out=os.path.join(vdir, f"libgsr_{name}.so"))
