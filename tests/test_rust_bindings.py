"""rust/heat-b200-sys/src/lib.rs cannot be compiled here (no Rust toolchain): check its extern "C" block against
include/heat_b200.h instead — every exported name must be bound, with the same number of arguments."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _strip(text, comment):
    return re.sub(comment, "", text, flags=re.S)


def header_prototypes():
    text = _strip(open(os.path.join(ROOT, "include", "heat_b200.h")).read(), r"/\*.*?\*/")
    out = {}
    for name, args in re.findall(r"\b(hb_[a-z0-9_]+)\s*\(([^;{}]*?)\)\s*;", text):
        args = args.strip()
        out[name] = 0 if args in ("", "void") else len(args.split(","))
    return out


def rust_prototypes():
    text = open(os.path.join(ROOT, "rust", "heat-b200-sys", "src", "lib.rs")).read()
    text = _strip(text, r"//[^\n]*")
    block = re.search(r'extern\s+"C"\s*\{(.*)\}', text, flags=re.S).group(1)
    out = {}
    for name, args in re.findall(r"pub\s+fn\s+(hb_[a-z0-9_]+)\s*\(([^)]*)\)", block):
        args = args.strip()
        out[name] = 0 if not args else len([a for a in args.split(",") if a.strip()])
    return out


def test_every_header_symbol_is_bound_with_the_same_arity():
    c, r = header_prototypes(), rust_prototypes()
    assert len(c) >= 30
    assert sorted(c) == sorted(r), (sorted(set(c) - set(r)), sorted(set(r) - set(c)))
    for name in c:
        assert c[name] == r[name], (name, c[name], r[name])


def test_struct_fields_match_the_header():
    """Field names (in order) of every #[repr(C)] struct against the C typedefs."""
    ctext = _strip(open(os.path.join(ROOT, "include", "heat_b200.h")).read(), r"/\*.*?\*/")
    rtext = _strip(open(os.path.join(ROOT, "rust", "heat-b200-sys", "src", "lib.rs")).read(), r"//[^\n]*")
    for struct in ("hb_layer", "hb_surface", "hb_batch_desc", "hb_model_info", "hb_surface_info", "hb_series",
                   "hb_ensemble_stats"):
        body = re.search(r"typedef struct \{([^}]*)\}\s*" + struct + r"\s*;", ctext, flags=re.S).group(1)
        cfields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            # "double a, b" / "const double* x" / "double normal[3]"
            names = decl.split(None, 1)[1] if not decl.startswith("const") else decl.split(None, 2)[2]
            for n in names.split(","):
                cfields.append(re.sub(r"[\*\s]|\[.*\]", "", n))
        rbody = re.search(r"pub struct " + struct + r"\s*\{([^}]*)\}", rtext, flags=re.S).group(1)
        rfields = re.findall(r"pub\s+([a-z0-9_]+)\s*:", rbody)
        assert cfields == rfields, (struct, cfields, rfields)
