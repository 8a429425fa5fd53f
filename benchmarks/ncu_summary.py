"""Cut an `ncu -i X.ncu-rep --page raw --csv` dump down to the columns the profiles/ summaries keep
(durations, DRAM bytes, pipe utilisation, occupancy, issue and stall counters).
Usage: python benchmarks/ncu_summary.py raw.csv > profiles/rN_x_ncu_full_summary.csv"""
import csv
import re
import sys

KEEP = re.compile(r"^(Kernel Name|Block Size|Grid Size)$|dram__bytes_(read|write)\.sum|gpu__dram_throughput|"
                  r"gpu__time_duration\.sum|launch__(occupancy|registers)|lts__t_sector_hit_rate|sm__cycles_(elapsed|active)|"
                  r"sm__inst_executed_pipe_(xu|fma|alu)|sm__pipe_tensor_cycles_active|sm__throughput|sm__warps_active|"
                  r"smsp__inst_executed\.sum|smsp__issue_active|smsp__pcsamp_warps_issue_stalled")


def main() -> None:
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = rows[0]
    cols = [i for i, name in enumerate(hdr) if KEEP.search(name)]
    out = csv.writer(sys.stdout)
    for r in rows:
        out.writerow([r[i] if i < len(r) else "" for i in cols])


if __name__ == "__main__":
    main()
