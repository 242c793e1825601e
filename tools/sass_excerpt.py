"""Writes the main-loop SASS of every tensor-core kernel (and of the fused all-reduce) to
profiles/sass/<kernel>.txt: the instructions around the tcgen05.mma issue (UTCHMMA), the TMA
loads (UTMALDG), the TMEM reads (LDTM) and, for the all-reduce, the peer / multimem accesses.

    python tools/sass_excerpt.py            # needs build/*.o (make -C csrc)

Complements tools/sass_summary.sh (mnemonic counts): these are listings a reader can follow.
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, 'profiles', 'sass')

# object file, regex on the demangled function name, output name, mnemonics that anchor excerpts
KERNELS = [
    ('conv_fprop.o', r'conv_fprop_tf32_kernel<256, 4, false, false>', 'fprop_bn256_1cta',
     ['UTCHMMA', 'UTMALDG', 'LDTM', 'UTMASTG']),
    ('conv_fprop.o', r'conv_fprop_tf32_pair_kernel<5>', 'fprop_bn256_cta_pair',
     ['UTCHMMA', 'UTMALDG', 'UTCBAR', 'LDTM']),
    ('conv_fprop.o', r'conv_fprop_tf32_kernel<64, 2, true, true>', 'fprop_bn64_halo_resident',
     ['UTCHMMA', 'UTMALDG']),
    ('conv_wgrad.o', r'conv_wgrad_tf32_kernel<256, 4>', 'wgrad_per_tap_bn256',
     ['UTCHMMA', 'UTMALDG', 'LDTM']),
    ('conv_wgrad_fused.o', r'conv_wgrad_fused_kernel<64>', 'wgrad_tap_fused_bn64',
     ['UTCHMMA', 'UTMALDG', 'LDTM']),
    ('allreduce.o', r'allreduce_scale_kernel<false, true>', 'allreduce_p2p_push',
     ['LDG.E.128', 'STG.E.128', 'LD.E.STRONG.SYS', 'ST.E.STRONG.SYS', 'NANOSLEEP']),
    ('allreduce.o', r'allreduce_scale_kernel<true, false>', 'allreduce_nvls',
     ['LDGMC', 'STG', 'RED', 'MULTIMEM']),
]


def functions(obj):
    """{demangled name: [sass lines]} of one object file."""
    text = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True).stdout
    out, name, body = {}, None, []
    for line in text.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            if name is not None:
                out[name] = body
            mangled = m.group(1)
            name = subprocess.run(['c++filt', mangled], capture_output=True,
                                  text=True).stdout.strip() or mangled
            body = []
        elif name is not None and re.match(r'\s+/\*[0-9a-f]{4,}\*/', line):
            body.append(re.sub(r'\s*/\* 0x[0-9a-f]+ \*/\s*$', '', line.rstrip()))
    if name is not None:
        out[name] = body
    return out


def excerpt(lines, anchors, context=14, limit=260):
    keep = set()
    for i, line in enumerate(lines):
        if any(a in line for a in anchors):
            keep.update(range(max(0, i - context), min(len(lines), i + context + 1)))
    out, last = [], None
    for i in sorted(keep):
        if last is not None and i != last + 1:
            out.append('        ...')
        out.append(lines[i])
        last = i
        if len(out) >= limit:
            out.append('        ... (truncated)')
            break
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    cache = {}
    for obj, pattern, name, anchors in KERNELS:
        path = os.path.join(ROOT, 'build', obj)
        if path not in cache:
            cache[path] = functions(path)
        match = [k for k in cache[path] if pattern in k]
        if not match:
            print(f'{name}: no function matching {pattern!r} in {obj}', file=sys.stderr)
            continue
        lines = cache[path][match[0]]
        counts = {a: sum(a in line for line in lines) for a in anchors}
        with open(os.path.join(OUT, name + '.txt'), 'w') as f:
            f.write(f'# {match[0][:160]}\n# from build/{obj} (nvcc 12.9, -gencode '
                    f'arch=compute_100a,code=sm_100a -O3 -lineinfo); {len(lines)} instructions\n'
                    f'# occurrences: {counts}\n# excerpt: +-14 instructions around each of '
                    f'{anchors}\n\n')
            f.write('\n'.join(excerpt(lines, anchors)) + '\n')
        print(name, len(lines), counts)


if __name__ == '__main__':
    main()
