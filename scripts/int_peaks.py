import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import package
tp = package()
names = {0:'IADD3',1:'IADD3+IMAD',2:'VIMNMX.U16x2',3:'VIADDMNMX.U16x2',4:'VIMNMX3.U16x2',5:'LDS.128 bcast + 4 VIADDMNMX'}
for v in range(6):
    print(v, names[v], '%.1f Gops/s (lane-ops)' % tp.measure_int_peak(0, v))
