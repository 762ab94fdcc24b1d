"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/cbx_t3.h
declares, and its host-only block pool matches the oracle bit-exactly.  No GPU compute."""
import json
import os
import re

import pytest

from chatterbox_vllm_b200 import _lib
from oracle import paging_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib.load()


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "cbx_t3.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(cbx_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"libcbx_t3.so does not export {name}"
    assert declared == set(_lib.EXPORTS)
    assert lib.cbx_abi_version() == 1


def test_block_pool_bit_exact_vs_oracle_and_vllm_golden(lib, golden_dir):
    from chatterbox_vllm_b200.ops import BlockPool
    g = json.load(open(os.path.join(golden_dir, "block_pool_vllm.json")))
    pool, ref = BlockPool(g["num_blocks"]), paging_ref.BlockPool(g["num_blocks"])
    held = {}
    for op, want in zip(g["script"], g["trace"]):
        if op[0] == "alloc":
            got = pool.alloc(op[2])
            assert got == want == ref.get_new_blocks(op[2])
            held.setdefault(op[1], []).extend(got)
        else:
            blks = list(reversed(held.pop(op[1])))
            pool.free(blks)
            ref.free_blocks(blks)
    assert pool.num_free() == g["free_after"] == ref.num_free()
    with pytest.raises(ValueError):
        pool.alloc(pool.num_free() + 1)
    assert pool.alloc(0) == []


def test_block_pool_random_script(lib):
    import random
    from chatterbox_vllm_b200.ops import BlockPool
    rng = random.Random(5)
    pool, ref, held = BlockPool(500), paging_ref.BlockPool(500), []
    for _ in range(400):
        if held and rng.random() < 0.45:
            blks = list(reversed(held.pop(rng.randrange(len(held)))))
            pool.free(blks), ref.free_blocks(blks)
        else:
            n = rng.randint(1, 40)
            if n <= ref.num_free():
                a, b = pool.alloc(n), ref.get_new_blocks(n)
                assert a == b
                held.append(a)
        assert pool.num_free() == ref.num_free()


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libcbx_t3.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()
