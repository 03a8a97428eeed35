#!/usr/bin/env python3
"""Symbolic per-basic-block expression printer for cuobjdump -sass output.

Used offline (no GPU) to read which mul/add pairs ptxas fused into FFMA in a
kernel, so that path-determining float formulas can be pinned with explicit
__fmaf_rn/__fmul_rn/__fadd_rn in our own kernels ("FMA profile").

usage: sass_expr.py file.sass function_substring [start_addr end_addr]
"""
import re, sys

INSN = re.compile(r'^\s*/\*([0-9a-f]{4,5})\*/\s+(?:(@!?U?P\d+)\s+)?([A-Z0-9_.]+)\s*(.*?)\s*;')

def load(path, fn):
    out, on = [], False
    for line in open(path):
        if 'Function :' in line:
            on = fn in line
            continue
        if not on:
            continue
        m = INSN.match(line)
        if m:
            out.append((int(m.group(1), 16), m.group(2), m.group(3), m.group(4)))
    return out

FP = {'FMUL', 'FFMA', 'FADD', 'MUFU', 'FSEL', 'FMNMX', 'FSETP', 'FCHK', 'I2F', 'I2FP', 'F2I', 'F2F', 'DSETP', 'FRND', 'SEL', 'MOV', 'IMAD', 'LD', 'LDG', 'LDC', 'LDL', 'LDS', 'HFMA2'}

def split_ops(s):
    ops, depth, cur = [], 0, ''
    for ch in s:
        if ch == '[': depth += 1
        if ch == ']': depth -= 1
        if ch == ',' and depth == 0:
            ops.append(cur.strip()); cur = ''
        else:
            cur += ch
    if cur.strip(): ops.append(cur.strip())
    return ops

def main():
    path, fn = sys.argv[1], sys.argv[2]
    lo = int(sys.argv[3], 16) if len(sys.argv) > 3 else 0
    hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else 1 << 30
    ins = load(path, fn)
    targets = set()
    for a, p, op, rest in ins:
        if op.startswith(('BRA', 'BSSY', 'CALL', 'BSYNC')):
            for t in re.findall(r'0x([0-9a-f]+)', rest):
                targets.add(int(t, 16))
    env = {}
    def val(tok, depth=0):
        m = re.match(r'^(-?)(\|?)(R\d+)(?:\.reuse)?(\|?)$', tok)
        if not m:
            return tok
        neg, ab, reg = m.group(1), m.group(2), m.group(3)
        e = env.get(reg, reg)
        if ab: e = 'abs(%s)' % e
        if neg: e = '-(%s)' % e
        return e
    for a, p, op, rest in ins:
        if a in targets:
            env = {}
            if lo <= a <= hi: print('--- block @%04x' % a)
        base = op.split('.')[0]
        ops = split_ops(rest)
        show = lo <= a <= hi
        if base in ('FMUL', 'FFMA', 'FADD', 'MUFU', 'FMNMX', 'FSEL', 'I2F', 'I2FP', 'FRND', 'F2F', 'F2I') and ops:
            dst = ops[0]
            args = [val(o) for o in ops[1:]]
            expr = '%s(%s)' % (op, ', '.join(args))
            if len(expr) > 400: expr = expr[:400] + '...'
            if show: print('%04x %s %s: t%04x = %s' % (a, p or '', dst, a, expr))
            if p is None and re.match(r'^R\d+$', dst):
                env[dst] = 't%04x' % a
            else:
                env.pop(dst, None)
        elif base in ('LD', 'LDG', 'LDL', 'LDS', 'LDC') and ops:
            dst = ops[0]
            m = re.search(r'\[(.*)\]', rest)
            addr = m.group(1) if m else '?'
            addr = re.sub(r'R\d+', lambda mm: mm.group(0), addr)
            wide = 2 if '.64' in op else 4 if '.128' in op else 1
            if show: print('%04x %s %s = %s[%s]' % (a, p or '', dst, op, addr))
            r0 = int(dst[1:]) if re.match(r'^R\d+$', dst) else None
            if r0 is not None:
                mo = re.search(r'\+0x([0-9a-f]+)', addr)
                off = int(mo.group(1), 16) if mo else 0
                for k in range(wide):
                    env['R%d' % (r0 + k)] = 'M[%s+%d]' % (re.sub(r'\.64.*', '', addr.split('+')[0]), off + 4 * k)
        elif base in ('FSETP', 'DSETP', 'FCHK'):
            args = [val(o) for o in ops]
            if show: print('%04x %s %s %s' % (a, p or '', op, ', '.join(args)))
        else:
            # any other writer kills the symbolic value of its destination
            if ops and re.match(r'^R\d+$', ops[0]):
                if base in ('MOV',) and len(ops) == 2 and p is None:
                    env[ops[0]] = val(ops[1])
                elif base == 'IMAD' and 'MOV' in op and p is None:
                    env[ops[0]] = val(ops[-1])
                else:
                    env.pop(ops[0], None)
                    if '.64' in op or 'WIDE' in op:
                        env.pop('R%d' % (int(ops[0][1:]) + 1), None)
            if show and base in ('BRA', 'CALL', 'EXIT', 'RET'):
                print('%04x %s %s %s' % (a, p or '', op, rest))

if __name__ == '__main__':
    main()
