"""bench.py and __graft_entry__.py only run end to end on a GPU box; a misspelt or deleted module-level name in them
would otherwise show up there first.  Every global a function of theirs loads must exist (CPU check, no CUDA)."""
import builtins
import dis
import importlib.util
import os
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name + "_static_check", os.path.join(ROOT, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)  # definitions only: the entry points sit behind `if __name__ == "__main__"`
    return mod


def _code_objects(code):
    yield code
    for c in code.co_consts:
        if isinstance(c, types.CodeType):
            yield from _code_objects(c)


@pytest.mark.parametrize("name", ["bench", "__graft_entry__"])
def test_every_global_a_function_loads_is_defined(name):
    mod = _load(name)
    missing = []
    for obj in vars(mod).values():
        funcs = [obj] if isinstance(obj, types.FunctionType) else \
            [m for m in vars(obj).values() if isinstance(m, types.FunctionType)] if isinstance(obj, type) else []
        for fn in funcs:
            if fn.__module__ != mod.__name__:
                continue
            for code in _code_objects(fn.__code__):
                for ins in dis.get_instructions(code):
                    if ins.opname in ("LOAD_GLOBAL", "LOAD_NAME") and ins.argval not in vars(mod) \
                            and not hasattr(builtins, ins.argval):
                        missing.append((fn.__name__, ins.argval))
    assert not missing, missing


def test_bench_line_constants():
    b = _load("bench")
    assert "Mrays/s" in b.METRIC and "1920x1080" in b.WORKLOAD and "64 spp" in b.WORKLOAD
    assert (b.WIDTH, b.HEIGHT, b.SPP, b.MAX_BOUNCES) == (1920, 1080, 64, 8)
    assert isinstance(b.NCU_TRAFFIC, int) and b.NCU_TRAFFIC > 0 and "ncu" in b.NCU_TRAFFIC_SOURCE
