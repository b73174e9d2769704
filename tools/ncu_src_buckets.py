"""Summarise the source page of an ncu report: warp-stall samples bucketed along the SASS of one kernel (dev tool).
  python tools/ncu_src_buckets.py report.ncu-rep [bucket]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 60
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = dict(zip(rows[0], rows[2]))
for k in ('Kernel Name', 'gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
          'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum'):
  print(k, d.get(k))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
isrc, iall, iex = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
data = [(int(r[iall]), int(r[iex]), r[isrc].strip()) for r in rows[2:] if len(r) > iall and r[iall].isdigit()]
if len(data) % 2 == 0 and [x[2] for x in data[:len(data) // 2]] == [x[2] for x in data[len(data) // 2:]]:
  data = data[len(data) // 2:]
tot = sum(x[0] for x in data)
print('samples', tot, 'instructions', len(data))
for s in range(0, len(data), bucket):
  seg = data[s:s + bucket]
  smp, ex = sum(x[0] for x in seg), sum(x[1] for x in seg)
  kinds = {}
  for x in seg:
    t = x[2].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    kinds[op] = kinds.get(op, 0) + 1
  top = sorted(kinds.items(), key=lambda kv: -kv[1])[:5]
  hot = max(seg, key=lambda x: x[0])
  if smp * 200 >= tot:
    print(f'{s:5d} {smp:6d} ({100 * smp / tot:4.1f}%) exec/line {ex // len(seg):8d} {top}  hot: {hot[0]} {hot[2][:60]}')
