"""CPU: the host core behind the C-ABI (no GPU calls): option grammar, dimension / type
strings, flexible-tensor header, dimension conversion.  The negative strings are the option
conformance suite of the reference (tests/nnstreamer_plugins/unittest_plugins.cc:290-950)."""
import ctypes as C

import numpy as np
import pytest

from nnstreamer_b200 import _lib as L

lib = L.lib()


def parse(mode, option):
    return L.parse_option(mode, option)


# ---- negative option strings (each must leave `option` unset in the reference) ---------
NEGATIVE = [
    ("dimchg", "10:11"), ("dimchg", "1,2"), ("dimchg", "0:2,1:3"),
    ("arithmetic", "typecast"), ("arithmetic", "typecast:unknown"), ("arithmetic", "typecast:char"),
    ("arithmetic", "typecast:double"), ("arithmetic", "add:0xF"), ("arithmetic", "add:1U"),
    ("arithmetic", "+2"), ("arithmetic", "*2"), ("arithmetic", "/2"), ("arithmetic", "max"),
    ("arithmetic", "per-channel:false"), ("arithmetic", "per-channel:invalid,add:1@2"),
    ("transpose", "5:2:4:3"), ("transpose", "2:3:1:0"), ("transpose", "0:3"),
    ("clamp", "50:20"), ("clamp", "+-1"), ("clamp", "+1:-2"), ("clamp", "1:1.7e309"),
    ("clamp", "-1.7e309:1"),
    ("stand", "invalid"), ("stand", "dc-average:unknown"), ("stand", "dc-average,per-channel"),
    ("stand", "dc-average:uint8,per-channel:yes"), ("stand", "dc-average:uint8,true"),
    # further strings derived from the grammars (gsttensor_transform.c:70-80)
    ("dimchg", "8:0"), ("dimchg", "0:8"), ("dimchg", "0:"), ("dimchg", ""), ("dimchg", " 0:1"),
    ("typecast", "unknown"), ("typecast", "float8"), ("typecast", " uint8"), ("typecast", "uint8 "),
    ("typecast", "int128"), ("typecast", "ufloat32"),
    ("arithmetic", "add"), ("arithmetic", "add:"), ("arithmetic", "add:1."), ("arithmetic", "add:1e"),
    ("arithmetic", "add:1,,mul:2"), ("arithmetic", "add:1@"), ("arithmetic", "sub:1"),
    ("arithmetic", "typecast:float32"), ("arithmetic", "add:1,typecast:uint8,mul:2"),
    ("arithmetic", "per-channel:true,add:1"), ("arithmetic", "per-channel:true@,add:1"),
    ("arithmetic", ",add:1"), ("arithmetic", "add:1 ,mul:2"), ("arithmetic", ""),
    ("transpose", "0:1:2"), ("transpose", "0:0:1:3"), ("transpose", "0:1:2:4"), ("transpose", "0:1:3:2"),
    ("transpose", "0:1:2:3:4"), ("transpose", "a:b:c:3"),
    ("clamp", "1"), ("clamp", "1:"), ("clamp", ":1"), ("clamp", "1:2:3"), ("clamp", "a:b"),
    ("clamp", "1e:2"), ("clamp", "1e-400:1"),
    ("stand", "default:"), ("stand", "default,"), ("stand", "per-channel:true"),
    ("stand", "default,per-channel:true,"), ("stand", "default:float32:uint8"),
]


@pytest.mark.parametrize("mode,option", NEGATIVE, ids=["%s|%s" % n for n in NEGATIVE])
def test_invalid_options_are_rejected(mode, option):
    code, opt = parse(mode, option)
    assert code == L.EINVAL
    assert opt.mode == -1
    assert lib.nnsb_last_error()


def test_dimchg_and_typecast_and_transpose_options():
    for f in range(8):
        for t in range(8):
            code, opt = parse("dimchg", "%d:%d" % (f, t))
            assert code == 0 and (opt.dimchg_from, opt.dimchg_to) == (f, t)
    for name in L.TYPES:
        for s in (name, name.upper(), name.capitalize()):
            code, opt = parse("typecast", s)
            assert code == 0 and opt.typecast_to == L.TYPE_ID[name], s
    import itertools
    for p in itertools.permutations((0, 1, 2)):
        code, opt = parse("transpose", "%d:%d:%d:3" % p)
        assert code == 0 and list(opt.trans_order)[:4] == list(p) + [3]


def _ops(opt):
    return [(opt.ops[i].op, opt.ops[i].applying_ch, opt.ops[i].is_float,
             opt.ops[i].dval if opt.ops[i].is_float else opt.ops[i].ival, opt.ops[i].cast_type)
            for i in range(opt.n_ops)]


def test_arithmetic_option_semantics():
    F32, END = L.TYPE_ID["float32"], L.TYPE_END
    code, opt = parse("arithmetic", "typecast:float32,add:-127.5,div:127.5")
    assert code == 0 and opt.arith_out_type == F32 and not opt.arith_per_channel
    assert _ops(opt) == [(0, -1, 0, 0, F32), (1, -1, 1, -127.5, END), (3, -1, 1, 127.5, END)]
    # literal class: '.', 'e', 'E' => double, else int64 (:796-807)
    code, opt = parse("arithmetic", "add:-10,mul:2.0,div:1e1,add:+3,mul:1E0")
    assert [o[2] for o in _ops(opt)] == [0, 1, 1, 0, 1]
    assert [o[3] for o in _ops(opt)] == [-10, 2.0, 10.0, 3, 1.0]
    # extra :operands are ignored (transform_arithmetic/runTest.sh:67-70)
    code, opt = parse("arithmetic", "add:9.9e-1:8.8:7")
    assert code == 0 and _ops(opt) == [(1, -1, 1, 9.9e-1, END)]
    # a typecast that is not first is removed (:728-737); here it is last -> trailing comma is fine
    code, opt = parse("arithmetic", "typecast:float64,add:0.2,add:0.1,typecast:uint16")
    assert code == 0 and opt.arith_out_type == L.TYPE_ID["float64"] and opt.n_ops == 3
    # per-channel (:760-773, :809-811)
    code, opt = parse("arithmetic", "typecast:int32,per-channel:true@0,mul:1@0,mul:2@1,add:-1@0,add:7")
    assert code == 0 and opt.arith_per_channel == 1 and opt.arith_ch_dim == 0
    assert [(o[0], o[1]) for o in _ops(opt)] == [(0, -1), (2, 0), (2, 1), (1, 0), (1, -1)]
    # @CH without per-channel:true is parsed but not honoured
    code, opt = parse("arithmetic", "per-channel:false,add:255@0")
    assert code == 0 and not opt.arith_per_channel and _ops(opt)[0][1] == -1
    code, opt = parse("arithmetic", "add:255@0")
    assert code == 0 and _ops(opt)[0][1] == -1
    # case-insensitive keywords; operators may follow each other without a comma (the regex's "(,|)")
    code, opt = parse("arithmetic", "TYPECAST:FLOAT32,ADD:1,Mul:2")
    assert code == 0 and opt.arith_out_type == F32 and opt.n_ops == 3
    code, opt = parse("arithmetic", "add:1mul:2")
    assert code == 0 and _ops(opt) == [(1, -1, 0, 1, END)]
    code, opt = parse("arithmetic", "add:1,")
    assert code == 0 and opt.n_ops == 1
    # @CH only read from the first operand (values = split (str_op[1], "@"))
    code, opt = parse("arithmetic", "per-channel:true@1,add:1:2@3")
    assert code == 0 and _ops(opt)[0][1] == -1
    code, opt = parse("arithmetic", "per-channel:true@12,add:1@3")
    assert code == 0 and opt.arith_ch_dim == 12 and _ops(opt)[0][1] == 3


def test_stand_and_clamp_options():
    code, opt = parse("stand", "default")
    assert code == 0 and (opt.stand_mode, opt.stand_out_type, opt.stand_per_channel) == (0, L.TYPE_END, 0)
    code, opt = parse("stand", "dc-average:float32,per-channel:true")
    assert code == 0 and (opt.stand_mode, opt.stand_out_type, opt.stand_per_channel) == (1, L.TYPE_ID["float32"], 1)
    code, opt = parse("stand", "DEFAULT:UINT8,PER-CHANNEL:FALSE")
    assert code == 0 and (opt.stand_mode, opt.stand_out_type, opt.stand_per_channel) == (0, L.TYPE_ID["uint8"], 0)
    for s, lo, hi in (("-50:50", -50.0, 50.0), ("+20:80", 20.0, 80.0), ("-33.3:+77.7", -33.3, 77.7),
                      (".5:1e3", 0.5, 1000.0), ("1:1", 1.0, 1.0), ("-1.5E-3:2e+2", -1.5e-3, 200.0)):
        code, opt = parse("clamp", s)
        assert code == 0 and (opt.clamp_min, opt.clamp_max) == (lo, hi), s


def test_type_and_dimension_helpers():
    for i, name in enumerate(L.TYPES):
        assert lib.nnsb_type_from_string(name.encode()) == i
        assert lib.nnsb_type_from_string(("  " + name.upper() + " ").encode()) == i
        assert lib.nnsb_type_to_string(i) == name.encode()
        assert lib.nnsb_element_size(i) == np.dtype(name).itemsize
    assert lib.nnsb_type_from_string(b"double") == L.TYPE_END
    assert lib.nnsb_type_from_string(b"") == L.TYPE_END
    assert lib.nnsb_type_to_string(L.TYPE_END) is None
    assert lib.nnsb_element_size(L.TYPE_END) == 0

    dim = (C.c_uint32 * 8)()
    # gst_tensor_parse_dimension: missing trailing dims = 1, blanks stripped, rank returned
    for s, rank, want in (("3:224:224:1", 4, [3, 224, 224, 1, 1, 1, 1, 1]),
                          ("5", 1, [5, 1, 1, 1, 1, 1, 1, 1]),
                          (" 3 : 4 ", 2, [3, 4, 1, 1, 1, 1, 1, 1]),
                          ("1:2:3:4:5:6:7:8", 8, [1, 2, 3, 4, 5, 6, 7, 8]),
                          ("3:4:", 2, [3, 4, 1, 1, 1, 1, 1, 1]),
                          ("", 0, [1] * 8)):
        assert lib.nnsb_dim_parse(s.encode(), dim) == rank, s
        assert list(dim) == want, s
    assert lib.nnsb_dim_parse(None, dim) == 0
    buf = C.create_string_buffer(128)
    d = (C.c_uint32 * 8)(3, 224, 224, 1, 1, 1, 1, 1)
    assert lib.nnsb_dim_to_string(d, 0, buf, 128) > 0 and buf.value == b"3:224:224:1:1:1:1:1"
    assert lib.nnsb_dim_to_string(d, 4, buf, 128) > 0 and buf.value == b"3:224:224:1"
    assert lib.nnsb_dim_to_string(d, 4, buf, 5) == L.EINVAL
    assert lib.nnsb_dim_is_valid(d) == 1
    d[7] = 0
    assert lib.nnsb_dim_is_valid(d) == 0
    d[7] = 1
    assert lib.nnsb_element_count(d) == 3 * 224 * 224
    info = L.make_info("float32", [3, 224, 224, 1])
    assert lib.nnsb_info_size(C.byref(info)) == 3 * 224 * 224 * 4


def test_flexible_header_round_trip_and_layout():
    info = L.make_info("float32", [3, 224, 224, 2])
    meta = L.MetaInfo()
    assert lib.nnsb_meta_from_info(C.byref(info), C.byref(meta)) == 0
    hdr = (C.c_uint8 * 128)(*([0xAA] * 128))
    assert lib.nnsb_meta_write_header(C.byref(meta), hdr) == 0
    words = np.frombuffer(bytes(hdr), dtype=np.uint32)
    # version 0xDE001000 (util_impl.c:1216-1226), type enum value, dims, format static, media tensor
    assert words[0] == 0xDE001000 and words[1] == L.TYPE_ID["float32"]
    assert list(words[2:10]) == [3, 224, 224, 2, 1, 1, 1, 1] and not words[10:18].any()
    assert words[18] == 0 and words[19] == 4 and not words[20:].any()
    meta2 = L.MetaInfo()
    assert lib.nnsb_meta_parse_header(C.byref(meta2), hdr) == 0
    assert lib.nnsb_meta_data_size(C.byref(meta2)) == 3 * 224 * 224 * 2 * 4
    info2 = L.TensorInfo()
    assert lib.nnsb_meta_to_info(C.byref(meta2), C.byref(info2)) == 0
    assert info2.type == info.type and list(info2.dim) == list(info.dim)
    bad = (C.c_uint8 * 128)()
    assert lib.nnsb_meta_parse_header(C.byref(meta2), bad) == L.EINVAL


def test_convert_dimension_both_directions():
    from oracle import oracle as O
    dim = [3, 10, 10, 4, 5, 6, 2, 7]
    info = L.make_info("uint8", dim)
    out = L.TensorInfo()
    for f in range(8):
        for t in range(8):
            code, opt = parse("dimchg", "%d:%d" % (f, t))
            assert lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SINK, C.byref(info), C.byref(out)) == 0
            want = O.convert_dims("dimchg", dim, from_=f, to=t)
            assert list(out.dim) == want, (f, t)
            # SRC direction is the inverse mapping
            back = L.TensorInfo()
            assert lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SRC, C.byref(out), C.byref(back)) == 0
            assert list(back.dim) == dim, (f, t)
    import itertools
    for p in itertools.permutations((0, 1, 2)):
        code, opt = parse("transpose", "%d:%d:%d:3" % p)
        lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SINK, C.byref(info), C.byref(out))
        assert list(out.dim) == O.convert_dims("transpose", dim, order=list(p) + [3])
        back = L.TensorInfo()
        lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SRC, C.byref(out), C.byref(back))
        assert list(back.dim) == dim
    # types: SINK direction sets the out type, SRC direction cannot know it (:1900-1947)
    for mode, option, sink_type in (("typecast", "float32", "float32"),
                                    ("arithmetic", "typecast:int16,add:1", "int16"),
                                    ("arithmetic", "add:1", "uint8"), ("stand", "default", "uint8"),
                                    ("stand", "dc-average:float64", "float64"), ("clamp", "0:1", "uint8")):
        code, opt = parse(mode, option)
        lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SINK, C.byref(info), C.byref(out))
        assert out.type == L.TYPE_ID[sink_type] and list(out.dim) == dim
        lib.nnsb_convert_dimension(C.byref(opt), L.PAD_SRC, C.byref(info), C.byref(out))
        unknowable = mode in ("typecast", "stand") or (mode == "arithmetic" and "typecast" in option)
        assert out.type == (L.TYPE_END if unknowable else info.type)
