"""Summarise .ncu-rep files (read here with `ncu -i ... --page raw --csv`) into a markdown table."""
import csv
import io
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "time us"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs"), ("launch__shared_mem_per_block_dynamic", "dyn smem KB"),
        ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %")]


def main():
    out = []
    for rep in sys.argv[1:]:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))
            name = d.get("Kernel Name", "?").replace("<unnamed>::", "")[:70]
            cells = []
            for k, _ in KEYS:
                v = d.get(k, "")
                try:
                    v = f"{float(v):.1f}" if "." in v else v
                except ValueError:
                    pass
                cells.append(f"{v} {u.get(k, '') if k.startswith('dram__bytes') else ''}".strip())
            out.append(f"| `{name}` | " + " | ".join(cells) + " |")
    print("| kernel | " + " | ".join(n for _, n in KEYS) + " |")
    print("|---|" + "---|" * len(KEYS))
    print("\n".join(out))


if __name__ == "__main__":
    main()
