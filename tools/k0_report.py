"""Prints the K0 DPX / integer-pipe micro-benchmark table (thread-instructions per second)."""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb
names = ["VIADDMNMX.S16x2", "VIMNMX3.S16x2", "VIADD.16x2", "gotoh-cell-mix(5)", "VIADDMNMX.S32", "IMAD", "cell-mix+3IMAD(8)", "HMNMX2 (fp16x2 max)", "HFMA2", "VIADDMNMX+HMNMX2 (2)", "VIADDMNMX+HFMA2 (2)", "VIADDMNMX+IMAD (2)",
         "VIADDMNMX imm addend", "VIADD.16x2 imm", "VIADDMNMX uniform addend", "VIADDMNMX imm max", "K1 cell imm penalties (5)", "K1 cell reg penalties (5)"]
out = {}
for k, n in enumerate(names):
    v = swb.dpx_peak(k)
    out[n] = v
    print("%-20s %8.3f T thread-instr/s  (%.1f lanes/clk/SM at 1965 MHz, 148 SMs)" % (n, v / 1e12, v / 148 / 1.965e9))
print(json.dumps(out))
