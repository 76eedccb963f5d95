"""ctypes loader for the reference's own compiled programs in oracle/_ref/ (CPU expansions, and the
reference's GPU expansions built for sm_100 as the on-box GPU comparison).

TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).  Needs no DaCe import:
the generated libraries export ``__dace_init_<n>(symbols...) -> state*``,
``__program_<n>(state*, arrays..., symbols...)`` and ``__dace_exit_<n>(state*)``
(reference dace/codegen/compiled_sdfg.py:233-240); argument order comes from the
manifest oracle/build_ref.py wrote when it compiled them.
"""
import ctypes
import json
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, '_ref')
MANIFEST = os.path.join(REF_DIR, 'manifest.json')

_CT = {'int64_t': ctypes.c_int64, 'int': ctypes.c_int, 'int32_t': ctypes.c_int32, 'long long': ctypes.c_longlong,
       'unsigned int': ctypes.c_uint, 'uint64_t': ctypes.c_uint64}


def available(ref_dir: str = REF_DIR) -> bool:
    return os.path.exists(os.path.join(ref_dir, 'manifest.json'))


def manifest(ref_dir: str = REF_DIR) -> dict:
    with open(os.path.join(ref_dir, 'manifest.json')) as f:
        return json.load(f)


class RefProgram:
    """One compiled reference SDFG.  Call with arrays (NumPy, C-contiguous) and symbols by name."""

    def __init__(self, name: str, ref_dir: str = REF_DIR):
        m = manifest(ref_dir)
        if name not in m['programs']:
            raise KeyError('no reference program %r in oracle/_ref (have: %s)' % (name, sorted(m['programs'])))
        self.name = name
        self.meta = m['programs'][name]
        self.lib = ctypes.CDLL(os.path.join(ref_dir, self.meta['lib']))
        self._init = getattr(self.lib, '__dace_init_' + name)
        self._init.restype = ctypes.c_void_p
        self._init.argtypes = [_CT[s['ctype']] for s in self.meta['init_symbols']]
        self._prog = getattr(self.lib, '__program_' + name)
        self._prog.restype = None
        self._prog.argtypes = [ctypes.c_void_p] + [
            ctypes.c_void_p if a['kind'] == 'array' else _CT[a['ctype']] for a in self.meta['args']
        ]
        self._exit = getattr(self.lib, '__dace_exit_' + name)
        self._exit.restype = ctypes.c_int
        self._exit.argtypes = [ctypes.c_void_p]
        self._state = None
        self._state_syms = None

    def __call__(self, **kwargs):
        syms = tuple(int(kwargs.get(s['name'], 0)) for s in self.meta['init_symbols'])
        if self._state is None or syms != self._state_syms:
            self.close()
            self._state = self._init(*syms)
            if not self._state:
                raise RuntimeError('__dace_init_%s failed (returned NULL)' % self.name)
            self._state_syms = syms
        call = []
        for a in self.meta['args']:
            v = kwargs[a['name']]
            if a['kind'] == 'array' and isinstance(v, int):
                call.append(v)   # raw (device) pointer for GPU_Global arrays of the reference's GPU programs
            elif a['kind'] == 'array':
                if not isinstance(v, np.ndarray) or not v.flags.c_contiguous:
                    raise ValueError('array argument %s must be a C-contiguous ndarray' % a['name'])
                call.append(v.ctypes.data)
            else:
                call.append(int(v))
        self._prog(self._state, *call)

    def close(self):
        if self._state is not None:
            self._exit(self._state)
            self._state = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
