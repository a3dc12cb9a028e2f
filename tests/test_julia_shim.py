"""CPU test of the Julia shim source (julia/QCB200.jl cannot be executed here: no Julia toolchain).
Every `ccall` in it must name a function include/qcb200.h declares, with the same number of
arguments and compatible scalar types; the Julia mirror structs must have the header's fields."""
import os
import re

from conftest import ROOT

JL = open(os.path.join(ROOT, "julia", "QCB200.jl")).read()
HDR = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "qcb200.h")).read(), flags=re.S)


def split_top(s):
    """Split on commas that are not inside brackets."""
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip()); cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def header_functions():
    fns = {}
    for m in re.finditer(r"\b([A-Za-z_][\w \*]*?)\b(qcb_[a-z_0-9]+)\s*\(([^;{]*?)\)\s*;", HDR, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        fns[name] = (ret, [] if args in ("", "void") else split_top(args))
    return fns


def julia_ccalls():
    """(name, return type, [argument types], number of call arguments) of every ccall."""
    calls = []
    for m in re.finditer(r"ccall\(\(:(qcb_[a-z_0-9]+), LIB\),\s*([\w\{\}]+),\s*\(", JL):
        i, depth = m.end(), 1
        while depth:                          # the tuple of argument types
            depth += {"(": 1, ")": -1}.get(JL[i], 0)
            i += 1
        types = split_top(JL[m.end():i - 1])
        j, depth = i, 1                       # the rest of the call: ", a, b, c)"
        while depth:
            depth += {"(": 1, ")": -1}.get(JL[j], 0)
            j += 1
        args = split_top(JL[i:j - 1].lstrip(", \n"))
        calls.append((m.group(1), m.group(2), types, len(args)))
    return calls


def c_class(t):
    t = t.replace("const ", "").strip()
    if "*" in t or "[" in t or t.endswith("handle") or t in ("qcb_handle", "qcb_fidelity", "qcb_aux", "qcb_free_phase", "qcb_problem"):
        return "ptr"
    t = t.split()[0]
    return {"int32_t": "i32", "int64_t": "i64", "double": "f64", "size_t": "u64", "int": "i32", "qcb_status": "i32"}.get(t, t)


def jl_class(t):
    if t.startswith(("Ptr", "Ref")) or t == "Cstring":
        return "ptr"
    return {"Int32": "i32", "Cint": "i32", "Int64": "i64", "Float64": "f64", "Csize_t": "u64", "Cvoid": "void"}.get(t, t)


def test_every_ccall_matches_the_header():
    fns = header_functions()
    calls = julia_ccalls()
    assert len(calls) >= 25
    for name, ret, types, nargs in calls:
        assert name in fns, f"{name} is not declared in include/qcb200.h"
        c_ret, c_args = fns[name]
        assert len(types) == len(c_args), f"{name}: {len(types)} argument types in the shim, {len(c_args)} in the header"
        assert nargs == len(c_args), f"{name}: {nargs} call arguments, {len(c_args)} parameters"
        for jt, ca in zip(types, c_args):
            # the last token of a C parameter is its name; arrays decay to pointers
            ctype = ca if ("*" in ca or "[" in ca) else " ".join(ca.split()[:-1])
            assert jl_class(jt) == c_class(ctype), f"{name}: Julia {jt} vs C '{ca}'"
        want = "void" if c_ret == "void" else c_class(c_ret)
        assert jl_class(ret) == want or (ret == "Cstring" and "char" in c_ret), f"{name}: return {ret} vs {c_ret}"


def test_shim_binds_the_hot_path_and_the_structure_queries():
    names = {c[0] for c in julia_ccalls()}
    for must in ("qcb_create", "qcb_destroy", "qcb_eval_defect", "qcb_eval_jacobian", "qcb_eval_hessian", "qcb_eval_all",
                 "qcb_set_structure", "qcb_jacobian_structure", "qcb_hessian_structure", "qcb_fidelity_eval",
                 "qcb_free_phase_eval", "qcb_aux_sensitivity_eval", "qcb_aux_regularizer_eval"):
        assert must in names, must
    for method in ("evaluate!", "jacobian_structure", "hessian_structure", "jacobian_values!", "hessian_of_lagrangian_values!",
                   "replace_integrators!", "DensityTrajectory", "MultiKetTrajectory", "SamplingTrajectory",
                   "VariationalKetIntegrator", "VariationalUnitaryIntegrator", "UnitarySmoothPulseProblem",
                   "QuantumStateSmoothPulseProblem", "UnitarySamplingProblem"):
        assert method in JL, method


def _struct_fields(src, name, julia):
    if julia:
        body = re.search(r"struct %s\b(.*?)\nend" % name, src, flags=re.S).group(1)
        body = re.sub(r"#.*", "", body)
        return re.findall(r"(\w+)::", body)
    body = re.search(r"typedef struct \{([^{}]*)\} %s;" % name, src).group(1)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        typ_stripped = re.sub(r"^(const\s+)?\w+\s*\**", "", decl, count=1)
        names += [re.sub(r"[\s\*]", "", x) for x in typ_stripped.split(",")]
    return names


def test_mirror_structs_have_the_header_fields():
    assert _struct_fields(JL, "QcbDesc", True) == _struct_fields(HDR, "qcb_desc", False)
    assert _struct_fields(JL, "QcbAuxDesc", True) == _struct_fields(HDR, "qcb_aux_desc", False)
    assert _struct_fields(JL, "QcbProblemDesc", True) == _struct_fields(HDR, "qcb_problem_desc", False)


def test_shim_binds_the_five_ipopt_callbacks():
    names = {c[0] for c in julia_ccalls()}
    for must in ("qcb_problem_create", "qcb_problem_destroy", "qcb_problem_dims", "qcb_problem_jacobian_structure",
                 "qcb_problem_hessian_structure", "qcb_problem_eval_f", "qcb_problem_eval_grad_f", "qcb_problem_eval_g",
                 "qcb_problem_eval_jac_g", "qcb_problem_eval_h", "qcb_problem_last_error"):
        assert must in names, must
