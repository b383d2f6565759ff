"""Four tcgen05 GEMM launches on training-step shapes, for an ncu capture (development probe)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'tools'))
import gemm_probe as gp
for s in gp.SHAPES:
    if s[0] in ('psa dX', 'psa dW', 'lin 64x456', 'wide dW', 'lin 456x64'):
        print(s[0], gp.run(*s, 3, reps=1))
