import sys; sys.path.insert(0, '.')
import open_spiel_1_b200 as bsg
e = bsg.Engine("connect_four", 0)
for k in range(7):
  print("kind", k, "%.2f T thread-instr/s" % (e.int_peak(k, 4000) / 1e12), flush=True)
