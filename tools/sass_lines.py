"""Per-source-line instruction / stall-sample totals of one kernel of an .ncu-rep, by joining ncu's SASS page (per-instruction
counters) with nvdisasm -g line info of the same cubin (instruction order is the same).
    python tools/sass_lines.py report.ncu-rep <launch index in the report> <mangled-name fragment, e.g. render_phaseILi1ELb0>
Needs the .so the report was taken with (simple_raytracer_b200/lib/libsrt_b200.so)."""
import csv, io, re, subprocess, sys, tempfile, os, collections

rep, idx, frag = sys.argv[1], int(sys.argv[2]), sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 45
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '--launch-skip', str(idx), '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
ins = [r for r in rows[2:] if len(r) == len(hdr)]
tmp = tempfile.mkdtemp()
lib = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'simple_raytracer_b200', 'lib', 'libsrt_b200.so')
subprocess.run(['cuobjdump', '-xelf', 'srt_phased', lib], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin') and f.startswith('srt_phased')][0]
dis = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
lines, cur, inside = [], None, False
for l in dis:
    if l.startswith('.text.'):
        inside = frag in l
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
        lines.append((cur, l.strip()))
print(f'{len(ins)} instructions in the report, {len(lines)} in the disassembly')
n = min(len(ins), len(lines))
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot = [0, 0, 0]
for k in range(n):
    r = ins[k]
    ie, te, sm = int(r[ix['Instructions Executed']]), int(r[ix['Thread Instructions Executed']]), int(r[ix['# Samples']])
    a = agg[lines[k][0]]
    a[0] += ie; a[1] += te; a[2] += sm; a[3] += 1
    tot[0] += ie; tot[1] += te; tot[2] += sm
print(f'total warp-instructions {tot[0]}, thread-instructions {tot[1]} ({tot[1] / max(tot[0], 1):.2f} per warp-instruction), samples {tot[2]}')
src = {}
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    if f not in src:
        try:
            src[f] = open(os.path.join(os.path.dirname(lib), '..', 'csrc', f)).read().splitlines()
        except OSError:
            src[f] = []
    text = src[f][ln - 1].strip()[:90] if 0 < ln <= len(src[f]) else ''
    print(f'{100 * a[0] / tot[0]:5.1f}% inst {100 * a[2] / max(tot[2], 1):5.1f}% samples  act {a[1] / max(a[0], 1):4.1f}  {f}:{ln}  [{a[3]} sass]  {text}')
