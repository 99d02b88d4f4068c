"""Start-up hook for ranks spawned by the dlbb launcher (bin/mpirun, bin/deepspeed, bin/b200run put this directory
first on PYTHONPATH, and Python imports ``sitecustomize`` from there before the user's script runs).

Its only job: make ``torch.distributed.init_process_group(backend='ccl')`` (and ``'dlbb'``) work in a script that
never imports ``oneccl_bindings_for_pytorch`` - the reference's test/test_deepseed.py:24 does exactly that and relies on
the oneCCL environment having registered the backend.  Nothing heavy happens here: no torch import, only an import hook
that, once ``torch.distributed`` has been loaded by the user's program, wraps ``init_process_group`` / ``new_group`` so
that the backend names are registered right before the first process group is created.

A ``sitecustomize`` that sits later on the path (the test harness of the GPU boxes installs one) is still executed.
"""
import importlib.abc
import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))


def _register_backends():
    try:
        from distributed_llm_backend_benchmark_b200.parallel.backend import register
    except Exception:   # the package is not importable from here: leave torch.distributed untouched
        return
    for name in ("ccl", "dlbb"):
        try:
            register(name)
        except Exception:
            pass


def _wrap(module):
    for fn_name in ("init_process_group", "new_group"):
        orig = getattr(module, fn_name, None)
        if orig is None or getattr(orig, "_dlbb_wrapped", False):
            continue

        def make(orig):
            def wrapper(*a, **kw):
                _register_backends()
                return orig(*a, **kw)
            wrapper.__name__ = orig.__name__
            wrapper.__doc__ = orig.__doc__
            wrapper.__wrapped__ = orig
            wrapper._dlbb_wrapped = True
            return wrapper

        setattr(module, fn_name, make(orig))


class _TorchDistributedHook(importlib.abc.MetaPathFinder):
    """Finds nothing itself; it only notices the import of torch.distributed and post-processes the module."""

    def find_spec(self, name, path, target=None):
        if name != "torch.distributed":
            return None
        sys.meta_path.remove(self)
        spec = importlib.util.find_spec(name)
        if spec is None or spec.loader is None or not hasattr(spec.loader, "exec_module"):
            return spec
        loader, orig_exec = spec.loader, spec.loader.exec_module

        def exec_module(module):
            orig_exec(module)
            try:
                _wrap(module)
            except Exception:
                pass

        try:
            loader.exec_module = exec_module
        except Exception:
            pass
        return spec


if os.environ.get("DLBB_NO_SITE_HOOK", "0") != "1":
    sys.meta_path.insert(0, _TorchDistributedHook())


def _chain_next_sitecustomize():
    for d in sys.path:
        try:
            if not d or os.path.abspath(d) == _HERE:
                continue
            cand = os.path.join(d, "sitecustomize.py")
            if os.path.isfile(cand):
                spec = importlib.util.spec_from_file_location("_dlbb_next_sitecustomize", cand)
                mod = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(mod)
                return
        except Exception:
            return


_chain_next_sitecustomize()
