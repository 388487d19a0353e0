"""Integer-pipe microbenchmarks (roofline denominators and pipe-sharing probes): python scripts/int_peaks.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
from __graft_entry__ import package
tp = package()
names = {0: "IADD3", 1: "IADD3+IMAD mix", 2: "VIMNMX.U16x2", 3: "VIADDMNMX.U16x2", 4: "VIMNMX3.U16x2", 5: "LDS.128 + 4 VIADDMNMX",
         6: "add-min steps: VIADDMNMX only", 7: "add-min steps: IMAD + VIMNMX3 only", 8: "add-min steps: half VIADDMNMX, half IMAD + VIMNMX3"}
for v in sorted(names):
    print("variant %d %-50s %8.1f G lane-results/s" % (v, names[v], tp.measure_int_peak(0, v)))
print("(variants 6-8 count 16 lane-results per 16 ADD-MIN STEPS of an accumulator set: compare them with each other)")
