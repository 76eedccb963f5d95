"""Every operator x element type on four layouts (512 MiB inputs): looks for operator / type pairs that miss a
streaming kernel (e.g. fall to the generic kernel), not for the last 10 %."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
BYTES = 1 << 29
OPS = {'sum': ('lambda a, b: a + b', 0), 'prod': ('lambda a, b: a * b', 1), 'max': ('lambda a, b: max(a, b)', None),
       'min': ('lambda a, b: min(a, b)', None), 'band': ('lambda a, b: a & b', -1), 'bor': ('lambda a, b: a | b', 0),
       'bxor': ('lambda a, b: a ^ b', 0), 'land': ('lambda a, b: a and b', 1), 'lor': ('lambda a, b: a or b', 0),
       'lxor': ('lambda a, b: a != b', 0)}
DTYPES = [torch.float32, torch.float64, torch.int32, torch.int64, torch.bfloat16, torch.float16, torch.uint8, torch.int8, torch.int16]
print('%-9s %-6s %s' % ('dtype', 'op', 'GB/s for: full | rows [.., 1024] | cols [.., 1024] | [256, 64, ..] axes (0, 2)'))
for dt in DTYPES:
    es = torch.empty((), dtype=dt).element_size()
    n = BYTES // es
    isf = dt.is_floating_point
    flat = torch.rand(n, device=dev).to(dt) if isf else torch.randint(0, 100, (n, ), device=dev, dtype=dt)
    layouts = [(flat, None), (flat.view(n // 1024, 1024), (1, )), (flat.view(n // 1024, 1024), (0, )),
               (flat.view(256, 64, n // (256 * 64)), (0, 2))]
    for op, (lam, ident) in OPS.items():
        if isf and op in ('band', 'bor', 'bxor', 'land', 'lor', 'lxor'):
            continue
        if ident is None:
            big = 60000 if es == 2 and isf else (3e38 if isf else (100 if es == 1 else 30000))
            ident = -big if op == 'max' else big
            if not isf and dt == torch.uint8 and op == 'max':
                ident = 0
        if ident == -1 and dt == torch.uint8:
            ident = 255
        cells = []
        for x, axes in layouts:
            try:
                out = db.reduce(lam, x, axis=axes, identity=ident)
                db.reduce(lam, x, out, axis=axes, identity=ident)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(3): db.reduce(lam, x, out, axis=axes, identity=ident)
                e1.record(); torch.cuda.synchronize()
                gbs = x.numel() * es / (e0.elapsed_time(e1) / 3) / 1e6
                cells.append('%6.0f%s' % (gbs, '!' if gbs < 3000 else ' '))
            except Exception as ex:
                cells.append(type(ex).__name__[:7])
        print('%-9s %-6s %s' % (str(dt).replace('torch.', ''), op, ' | '.join(cells)), flush=True)
    del flat, layouts
