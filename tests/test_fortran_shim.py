"""The ISO_C_BINDING shim cannot be compiled here (no Fortran compiler), so its interface blocks are held against the header statically: every
bind(C) name exists in include/ufe_b200.h with the same number of arguments, every entry of the path has a binding (only tooling is left out), and the
members of the config struct appear in the header's order."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# measurement / introspection helpers of bench.py and the tests; a Fortran host has no use for them
TOOLING = {"ufe_b200_get_mg_info", "ufe_b200_get_stiffness", "ufe_b200_get_stiffness_size", "ufe_b200_get_stiffness_storage_bytes", "ufe_b200_host_register",
           "ufe_b200_host_unregister", "ufe_b200_l2_flush", "ufe_b200_mg_host_transfer", "ufe_b200_mg_host_transfer_get", "ufe_b200_mg_host_transfer_parts",
           "ufe_b200_mg_host_transfer_parts_get", "ufe_b200_timer_start", "ufe_b200_timer_stop", "ufe_b200_timing_enable", "ufe_b200_timing_get",
           "ufe_b200_timing_reset"}


def header_prototypes():
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "ufe_b200.h")).read(), flags=re.S)
    out = {}
    for m in re.finditer(r"\b(?:int|void|long long|const char\*|double)\s+(ufe_b200_[A-Za-z0-9_]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out, hdr


def shim_bindings():
    src = open(os.path.join(ROOT, "ufemism2.0_b200", "fortran", "ufe_b200_mod.f90")).read()
    src = "\n".join(l.split("!")[0] for l in src.splitlines())            # comments off (no '!' occurs inside the strings of the interface part)
    src = re.sub(r"&\s*\n\s*", "", src)                                      # continuation lines joined
    out = {}
    for m in re.finditer(r"(?:function|subroutine)\s+(\w+)\s*\(([^)]*)\)\s*bind\(C,\s*name='(ufe_b200_\w+)'\)", src):
        assert m.group(1) == "c_" + m.group(3), m.group(1)
        args = m.group(2).strip()
        out[m.group(3)] = 0 if not args else len(args.split(","))
    return out, src


def test_every_binding_matches_a_header_prototype():
    protos, _ = header_prototypes()
    shim, _ = shim_bindings()
    assert len(protos) > 60 and len(shim) > 45
    assert not set(shim) - set(protos), set(shim) - set(protos)
    wrong = {k: (protos[k], shim[k]) for k in shim if protos[k] != shim[k]}
    assert not wrong, wrong
    assert set(protos) - set(shim) == TOOLING, (set(protos) - set(shim)) ^ TOOLING


def test_config_struct_members_follow_the_header_order():
    _, hdr = header_prototypes()
    body = re.search(r"typedef struct ufe_b200_config \{(.*?)\} ufe_b200_config;", hdr, flags=re.S).group(1)
    c_members = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            c_members += [re.sub(r"\[.*?\]", "", n).strip() for n in re.sub(r"^(double|int32_t)\s+", "", decl).split(",")]
    _, src = shim_bindings()
    fbody = re.search(r"type, bind\(C\) :: ufe_b200_config(.*?)end type", src, flags=re.S).group(1)
    f_members = []
    for line in fbody.splitlines():
        if "::" in line:
            f_members += [re.sub(r"\(.*?\)", "", n).strip() for n in line.split("::")[1].split(",")]
    assert f_members == c_members, [(a, b) for a, b in zip(f_members, c_members) if a != b][:5]


def test_ctypes_config_follows_the_header_order_and_size():
    import ctypes as C
    import ufe_b200 as ub
    _, hdr = header_prototypes()
    body = re.search(r"typedef struct ufe_b200_config \{(.*?)\} ufe_b200_config;", hdr, flags=re.S).group(1)
    c_members, size = [], 0
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        width = 8 if decl.startswith("double") else 4
        for n in re.sub(r"^(double|int32_t)\s+", "", decl).split(","):
            dim = re.search(r"\[(\d+)\]", n)
            c_members.append(re.sub(r"\[.*?\]", "", n).strip())
            size += width * (int(dim.group(1)) if dim else 1)
    assert [n for n, _ in ub.Config._fields_] == c_members
    assert C.sizeof(ub.Config) == (size + 7) // 8 * 8            # doubles first, then 32-bit integers: no padding except at the end


def _struct_members(hdr, name):
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), hdr, flags=re.S).group(1)
    out = []
    for decl in body.split(";"):
        decl = re.sub(r"^(const\s+)?(double|int32_t)\s*\*?", "", decl.strip())
        out += [n.replace("*", "").strip() for n in decl.split(",") if n.strip()]
    return out


def test_pointer_tables_of_the_wrappers_follow_the_struct_member_order():
    """The Fortran wrappers hand structs of pointers over as arrays of c_loc's: their order must be the member order of the C struct."""
    import ufe_b200 as ub
    _, hdr = header_prototypes()
    _, src = shim_bindings()
    tables = {"ice_in": "ufe_b200_ice_inputs", "state": "ufe_b200_velocity_state", "out": "ufe_b200_ice_velocities", "io": "ufe_b200_thickness_io",
              "ice_in_ssa": "ufe_b200_ice_inputs", "state_ssa": "ufe_b200_velocity_state"}
    rename = {"basal_friction_coefficient": "basal_friction_coefficient_a",            # ice%basal_friction_coefficient is the a-grid field
              "noice": "mask_noice", "bm": "BC_prescr_mask", "bh": "BC_prescr_Hi"}       # 0/1 copy of the logical mask; optional arguments
    ssa = {"u_b": "u_vav_b", "v_b": "v_vav_b", "eta_a": "eta_vav_a"}                     # how the SSA solver object's members travel
    for var, struct in tables.items():
        m = re.search(r"\n\s*%s = \[(.*?)\]" % var, src, flags=re.S)
        names = [re.search(r"%?(\w+)\s*\)?$", item.strip()).group(1) for item in m.group(1).split(",")]
        names = [rename.get(n, n) for n in names]
        members = _struct_members(hdr, struct)
        if var == "state_ssa":      # members the SSA has no field for are c_null_ptr; the others sit at their member's position
            assert len(names) == len(members) and names.count("c_null_ptr") == 16
            names = [m_ if n == "c_null_ptr" else ssa.get(n, n) for n, m_ in zip(names, members)]
        assert names == members, (var, names, members)
    assert _struct_members(hdr, "ufe_b200_velocity_state") == ub.VELOCITY_FIELDS and _struct_members(hdr, "ufe_b200_ice_velocities") == ub.ICE_VELOCITY_FIELDS
