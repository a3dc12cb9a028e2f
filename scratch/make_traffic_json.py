"""profiles/r01_traffic_bgemm_iso.json from an `ncu --set full` report: per-launch DRAM bytes, tensor
flops and DMMA-pipe activity of the bgemm_iso_kernel launches with a full per-drive grid."""
import csv, json, subprocess, sys
rep, src = sys.argv[1], sys.argv[2]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines())); h, units = rows[0], rows[1]; ix = {k: i for i, k in enumerate(h)}
def val(r, k):
    v = float(r[ix[k]].replace(',', '')); u = units[ix[k]]
    return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'us': 1.0, 'ms': 1e3, 'ns': 1e-3}.get(u, 1.0)
out = []
for r in rows[2:]:
    if 'bgemm_iso' not in r[ix['Kernel Name']]: continue
    grid = int(r[ix['launch__grid_size']].replace(',', ''))
    fl = val(r, 'sm__ops_path_tensor_src_fp64.sum'); t = val(r, 'gpu__time_duration.sum')
    out.append({'grid': grid, 'duration_us': t, 'dram_read_bytes': val(r, 'dram__bytes_read.sum'),
                'dram_write_bytes': val(r, 'dram__bytes_write.sum'), 'tensor_fp64_flops': fl,
                'dmma_pipe_active_pct': val(r, 'sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active'),
                'tflops': fl / t / 1e6})
big = max(o['grid'] for o in out)
pd = [o for o in out if o['grid'] == big]
ratio = sum(o['dram_read_bytes'] + o['dram_write_bytes'] for o in pd) / sum(o['tensor_fp64_flops'] for o in pd)
json.dump({'kernel': 'bgemm_iso_kernel<3>', 'source': src, 'launches': out,
           'dram_bytes_per_flop_per_drive_launches': ratio}, open('profiles/r01_traffic_bgemm_iso.json', 'w'), indent=1)
print(len(out), 'launches;', len(pd), 'per-drive; bytes/flop', ratio)
