"""CPU test: the ctypes mirrors in ggml_b200.py have the size the library was compiled with (b200_abi_sizeof, include/ggml-b200.h) --
a field added on one side only would otherwise shift every pointer behind it without any error at the call."""
import ctypes as C


def test_ctypes_mirrors_match_the_compiled_structs():
    import ggml_b200 as B
    lib = B.lib()
    for which, mirror in B.abi_mirrors().items():
        want = lib.b200_abi_sizeof(which)
        assert want > 0, which
        assert C.sizeof(mirror) == want, f"{mirror.__name__}: ctypes {C.sizeof(mirror)} bytes, library {want}"
    assert lib.b200_abi_sizeof(99) == 0


def test_chain_split_field_offsets():
    """b200_chain_split: {int32 world, rank; void * ll[8]; uint64 ll_bytes; int64 row_lo[6][3]}"""
    import ggml_b200 as B
    S = B.ChainSplit
    assert (S.world.offset, S.rank.offset, S.ll.offset, S.ll_bytes.offset, S.row_lo.offset) == (0, 4, 8, 72, 80)
    sp = S()
    sp.row_lo[2][1] = 7
    flat = (C.c_int64 * 18).from_buffer(sp, S.row_lo.offset)
    assert flat[2 * 3 + 1] == 7                                  # row-major [phase][result], as the C array
