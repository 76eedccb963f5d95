"""Minimal ctypes loader for DaCe-generated shared objects whose library nodes were expanded with
the ``'B200'`` implementations - what ``CompiledSDFG`` does on the reference side
(reference dace/codegen/compiled_sdfg.py:233-240 entry points, :549-592 ``fast_call``), without
importing DaCe, so a program built once (where the reference toolchain lives) runs on a box that
only has this package:

    ``__dace_init_<name>(symbols...) -> state*``
    ``__program_<name>(state*, arrays..., symbols...)``     arrays first (alphabetical), then symbols
    ``__dace_exit_<name>(state*) -> int``

The argument order comes from the ``manifest.json`` written next to the ``.so`` by the build script
(tests/dace_e2e/build_programs.py).  Arrays may be NumPy arrays, torch tensors (host or CUDA) or raw
integer pointers; nothing is copied or converted here.
"""
import ctypes
import json
import os

_CT = {'int64_t': ctypes.c_int64, 'int': ctypes.c_int, 'int32_t': ctypes.c_int32, 'long long': ctypes.c_longlong,
       'unsigned int': ctypes.c_uint, 'uint64_t': ctypes.c_uint64}


def manifest(directory: str) -> dict:
    with open(os.path.join(directory, 'manifest.json')) as f:
        return json.load(f)


def available(directory: str, name: str = None) -> bool:
    try:
        m = manifest(directory)
    except Exception:
        return False
    if name is None:
        return True
    return name in m['programs'] and os.path.exists(os.path.join(directory, m['programs'][name]['lib']))


def _pointer(v) -> int:
    if isinstance(v, int):
        return v
    if hasattr(v, 'data_ptr'):       # torch tensor, host (pinned) or device
        return int(v.data_ptr())
    if hasattr(v, 'ctypes'):         # NumPy array
        if not v.flags.c_contiguous:
            raise ValueError('array arguments must be C-contiguous')
        return int(v.ctypes.data)
    raise TypeError('cannot take the address of %r' % type(v))


class CompiledProgram:
    """One DaCe-built program.  ``prog(A=..., B=..., N=...)`` = ``CompiledSDFG.__call__``: the state is
    (re)initialised when the symbol values change, the call returns when the generated code returns
    (it synchronises its streams itself)."""

    def __init__(self, directory: str, name: str):
        m = manifest(directory)
        if name not in m['programs']:
            raise KeyError('no program %r in %s (have: %s)' % (name, directory, sorted(m['programs'])))
        self.name = name
        self.meta = m['programs'][name]
        self.path = os.path.join(directory, self.meta['lib'])
        self.lib = ctypes.CDLL(self.path)
        self._init = getattr(self.lib, '__dace_init_' + name)
        self._init.restype = ctypes.c_void_p
        self._init.argtypes = [_CT[s['ctype']] for s in self.meta['init_symbols']]
        self._prog = getattr(self.lib, '__program_' + name)
        self._prog.restype = None
        self._prog.argtypes = [ctypes.c_void_p] + [
            ctypes.c_void_p if a['kind'] == 'array' else _CT[a['ctype']] for a in self.meta['args']]
        self._exit = getattr(self.lib, '__dace_exit_' + name)
        self._exit.restype = ctypes.c_int
        self._exit.argtypes = [ctypes.c_void_p]
        self._state = None
        self._syms = None

    def bind(self, **kwargs):
        """Marshal once, call many times (the reference's ``fast_call``): returns a zero-argument callable."""
        syms = tuple(int(kwargs.get(s['name'], 0)) for s in self.meta['init_symbols'])
        if self._state is None or syms != self._syms:
            self.close()
            self._state = self._init(*syms)
            if not self._state:
                raise RuntimeError('__dace_init_%s failed' % self.name)
            self._syms = syms
        args = [_pointer(kwargs[a['name']]) if a['kind'] == 'array' else int(kwargs[a['name']])
                for a in self.meta['args']]
        prog, state = self._prog, self._state
        return lambda: prog(state, *args)

    def __call__(self, **kwargs):
        self.bind(**kwargs)()

    def close(self) -> int:
        rc = 0
        if self._state is not None:
            rc = int(self._exit(self._state))
            self._state = None
        return rc

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
