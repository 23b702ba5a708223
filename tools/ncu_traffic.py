#!/usr/bin/env python
"""Per-kernel DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum per launch, averaged over the launches of one
step) from the raw CSV of an `ncu --set full` capture of one training step -> the JSON bench.py reads for
`roofline.traffic` and the per-stage `dram_frac`.

    python tools/ncu_traffic.py profiles/r01_v8_step_full_raw.csv.gz profiles/r01_v8_traffic.json
"""
import csv
import gzip
import json
import sys

# kernel name prefix -> ABI stage it belongs to (the stage's other kernels are small)
STAGE = [('void tile_rows_kernel<0>', 'cn_project_gather_fwd (tile_rows)'), ('mlp_fwd_kernel<0', 'cn_mlp_feature_fwd'),
         ('mlp_fwd_kernel<1', 'cn_mlp_fusion_fwd'), ('mlp_block_bwd_kernel', 'cn_mlp_block_bwd'),
         ('mlp_l0_wgrad_kernel', 'cn_mlp_l0_wgrad'), ('mlp_l0_dgrad_kernel', 'cn_mlp_l0_dgrad'),
         ('void tile_rows_kernel<1>', 'cn_project_gather_bwd (tile_rows)'), ('feature_grad_kernel', 'cn_feature_grad'),
         ('void tile_rows_lr_kernel<0>', 'cn_project_gather_fwd (tile_rows)'),
         ('void tile_rows_lr_kernel<1>', 'cn_project_gather_bwd (tile_rows)'), ('feature_grad_lr_kernel', 'cn_feature_grad'),
         ('mlp_stage_bwd_kernel', 'cn_mlp_stage_bwd')]


def main(src, dst):
    op = gzip.open if src.endswith('.gz') else open
    rows = list(csv.reader(op(src, 'rt')))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def gb(r, name):
        v, u = float(r[col[name]]), units[col[name]]
        return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}[u]
    acc = {}
    total, total_us, n_all = 0.0, 0.0, 0
    for r in rows[2:]:
        name = r[col['Kernel Name']]
        total += gb(r, 'dram__bytes_read.sum') + gb(r, 'dram__bytes_write.sum')
        total_us += float(r[col['gpu__time_duration.sum']]) * {'us': 1.0, 'ms': 1e3, 'ns': 1e-3}[units[col['gpu__time_duration.sum']]]
        n_all += 1
        for prefix, stage in STAGE:
            if prefix.replace('void ', '') in name:
                a = acc.setdefault(stage, [0, 0.0, 0.0])
                a[0] += 1
                a[1] += gb(r, 'dram__bytes_read.sum') + gb(r, 'dram__bytes_write.sum')
                a[2] += float(r[col['gpu__time_duration.sum']]) * {'us': 1.0, 'ms': 1e3, 'ns': 1e-3}[units[col['gpu__time_duration.sum']]]
                break
    out = {k: {'launches_per_step': n, 'dram_bytes_per_launch': int(b / n), 'ncu_us_per_launch': round(t / n, 1)}
           for k, (n, b, t) in acc.items()}
    out['_step'] = {'launches': n_all, 'dram_bytes': int(total), 'ncu_us': round(total_us, 1)}
    out['_source'] = (f'{src}: ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum, one '
                      f'training step of `bench.py --steps 1 --warmup 3` (cfg2), averaged over the launches of each kernel')
    with open(dst, 'w') as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2])
