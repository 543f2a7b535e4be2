"""profiles/traffic.json from an `ncu --set full` capture of bench.py's kernel (the file bench.py
reads for roofline.traffic).  Run here (no GPU) on the .ncu-rep that scripts/gpu_ncu_bench.sh
brought back:

    python scripts/ncu_traffic.py gpurun_out/prof_bench.ncu-rep profiles/r02X_ncu_full_bench.csv [f64|f32]

Writes the key-metric summary CSV (scripts/ncu_summary.py format) and profiles/traffic.json with the
DRAM bytes of the ONE captured launch, its L2 hit rate, the kernel name and the git head."""
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
rep, out_csv = sys.argv[1], sys.argv[2]
dtype = sys.argv[3] if len(sys.argv) > 3 else 'f64'
subprocess.run([sys.executable, str(ROOT / 'scripts' / 'ncu_summary.py'), rep, out_csv], check=True,
               stdout=subprocess.DEVNULL)
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
d = dict(zip(rows[0], zip(rows[1], rows[2])))


def val(k):
    u, v = d[k]
    f = float(v.replace(',', ''))
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}.get(u, 1)
    return f * scale


rd, wr = val('dram__bytes_read.sum'), val('dram__bytes_write.sum')
head = subprocess.run(['git', 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True,
                      cwd=ROOT).stdout.strip()
def per_cycle(op):
    k = f'smsp__sass_thread_inst_executed_op_{op}_pred_on.sum.per_cycle_elapsed'
    return float(d[k][1].replace(',', '')) if k in d else 0.0


pre = 'f' if dtype == 'f32' else 'd'
cycles = float(d['sm__cycles_elapsed.max'][1].replace(',', '')) if 'sm__cycles_elapsed.max' in d else 0.0
flops_exec = (2.0 * per_cycle(pre + 'fma') + per_cycle(pre + 'add') + per_cycle(pre + 'mul')) * cycles
out = {
    'flops_executed_per_launch': flops_exec,   # thread-level FMA x 2 + ADD + MUL (predicated-on), whole launch
    'dram_bytes_per_launch': rd + wr, 'dram_bytes_read': rd, 'dram_bytes_write': wr,
    'lts_sector_hit_rate_pct': float(d['lts__t_sector_hit_rate.pct'][1]),
    'gpu_time_us': float(d['gpu__time_duration.sum'][1].replace(',', '')) *
    {'us': 1, 'ms': 1e3, 'ns': 1e-3, 's': 1e6}.get(d['gpu__time_duration.sum'][0], 1),
    'kernel': d['Kernel Name'][1], 'dtype': dtype,
    'launch': 'one timed step of `python bench.py --steps 3 --warmup 3 --no-cpu --no-extras` '
              '(4096 problems, cold start, SQP to convergence), captured by scripts/gpu_ncu_bench.sh',
    'source': f'{Path(out_csv).as_posix()} (ncu --set full --clock-control none, one launch)',
    'head': head,
}
(ROOT / 'profiles' / 'traffic.json').write_text(json.dumps(out, indent=1) + '\n')
print(json.dumps(out, indent=1))
