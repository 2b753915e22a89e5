import sys, os
sys.path.insert(0, os.getcwd())
import towr2025_b200 as T
s=T.make_spec(T.Biped,0,2.4); dt=0.025
s.dt_constraint_dynamic=dt; s.dt_constraint_range_of_motion=0.8*dt; s.dt_constraint_force=0.2*dt
st=T.Structure(s)
b=T.Batch(st, 64, 0)
print("ok")
