"""XT-32 scan leg of bench.py alone (A/B of library builds through LSB200_LIB): prints resident and end-to-end microseconds per scan."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import xt32_scan_leg
d = xt32_scan_leg(0, reps=2000, cpu=False)
print(os.environ.get("LSB200_LIB", "default")[-16:], round(d["resident_us_per_scan"], 2), round(d["e2e_us_per_scan"], 2))
