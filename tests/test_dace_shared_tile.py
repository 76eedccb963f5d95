"""CPU: the shared-tile plug-in (dace_b200/shared_tile.py) on the reference's code generator -
AccumulateTransient into shared memory keeps the reference's generated call sites, the .cu gains the device
header + a static_assert that its specialisations are the ones in use, and the header really specialises
DaCe's runtime templates (nvcc compile-only against the reference's runtime headers)."""
import os
import subprocess
import tempfile

import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='needs the reference checkout (DaCe frontend + codegen)')

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_accumulate_transient_in_shared_memory_uses_the_b200_header():
    dace = import_reference()
    from dace_b200 import shared_tile as ST
    N = dace.symbol('N', dace.int64)

    @dace.program
    def tile_hist(keys: dace.int32[N], bins: dace.int32[256]):
        for b in dace.map[0:N // 1024]:
            for t in dace.map[0:256]:
                for j in range(4):
                    with dace.tasklet:
                        k << keys[b * 1024 + j * 256 + t]
                        out >> bins(1, lambda x, y: x + y)[:]
                        out[k] = 1

    sdfg = tile_hist.to_sdfg(simplify=True)
    sdfg.name = 'b200_tile_codegen'
    sdfg.apply_gpu_transformations()
    assert ST.use_b200_shared_tiles(sdfg) == 0                      # nothing privatised yet
    st = next(s for s in sdfg.states() if sum(isinstance(n, dace.nodes.MapExit) for n in s.nodes()) >= 2)
    exits = [n for n in st.nodes() if isinstance(n, dace.nodes.MapExit)]
    inner = next(e for e in exits if e.map.params == ['t'])
    outer = next(e for e in exits if e.map.params == ['b'])
    st.entry_node(inner).map.schedule = dace.ScheduleType.GPU_ThreadBlock
    tile = ST.privatise_wcr_in_shared(sdfg, st, inner, outer)
    assert tile.desc(sdfg).storage == dace.StorageType.GPU_Shared and tile.setzero
    assert len(ST.shared_wcr_tiles(sdfg)) == 1
    cu = next(c.clean_code for c in sdfg.generate_code() if c.language == 'cu')
    # the reference's call sites, untouched ...
    assert 'dace::ResetShared<int, 256, 1, 1, 256, 1, false>::Reset(' in cu
    assert 'dace::SharedToGlobal1D<int, 256, 1, 1, 256, false>::template Accum<dace::ReductionType::Sum>(' in cu
    # ... served by the device header, which comes after <dace/dace.h> and before the kernel
    inc = cu.index('dace_b200_shared_accum.cuh')
    assert cu.index('#include <dace/dace.h>') < inc < cu.index('__global__')
    assert 'static_assert(dace::SharedToGlobal1D<int, 32, 1, 1, 32, false>::kDaceB200 == 1' in cu
    assert ST.use_b200_shared_tiles(sdfg) == 1                      # idempotent: nothing is added twice
    cu2 = next(c.clean_code for c in sdfg.generate_code() if c.language == 'cu')
    assert cu2.count('#include "') == cu.count('#include "') and cu2.count('static_assert(') == cu.count('static_assert(')


def test_header_specialises_the_runtime_templates():
    """nvcc -c of a translation unit shaped like a generated .cu: <dace/dace.h>, the header, static_asserts and a
    kernel with the generated call sites for every specialised element type."""
    from oracle.refenv import REFERENCE_ROOT
    rt = os.path.join(REFERENCE_ROOT, 'dace', 'runtime', 'include')
    if not os.path.exists(os.path.join(rt, 'dace', 'dace.h')):
        pytest.skip('reference runtime headers not found')
    types = ['float', 'double', 'int', 'unsigned int', 'long long', 'unsigned long long']
    body = ['#include <cuda_runtime.h>', '#include <dace/dace.h>', '#include "dace_b200_shared_accum.cuh"']
    for i, t in enumerate(types):
        body.append('static_assert(dace::SharedToGlobal1D<%s, 128, 2, 1, 512, false>::kDaceB200 == 1, "");' % t)
        body.append('static_assert(dace::ResetShared<%s, 128, 2, 1, 512, 1, false>::kDaceB200 == 1, "");' % t)
        body.append('''__global__ void k%d(%s* g) {
    __shared__ %s tile[512];
    dace::ResetShared<%s, 128, 2, 1, 512, 1, false>::Reset(tile);
    dace::SharedToGlobal1D<%s, 128, 2, 1, 512, false>::template Accum<dace::ReductionType::Sum>(tile, 1, g, 1);
    dace::SharedToGlobal1D<%s, 128, 2, 1, 512, true>::template Accum<dace::ReductionType::Max>(tile, 1, g, 2);
    dace::SharedToGlobal1D<%s, 128, 2, 1, 512, true>::Copy(tile, 1, g, 1);
}''' % (i, t, t, t, t, t, t))
    # a type without a specialisation keeps the reference's implementation
    body.append('struct pod { int a; };')
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, 'tu.cu')
        open(src, 'w').write('\n'.join(body) + '\n')
        res = subprocess.run(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-std=c++17', '-c', src, '-o',
                              os.path.join(tmp, 'tu.o'), '-I', rt, '-I', os.path.join(ROOT, 'include'),
                              '-ccbin', '/usr/bin/g++'],
                             capture_output=True, text=True)
        assert res.returncode == 0, res.stderr[-3000:]
