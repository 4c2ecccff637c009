"""FastDiv::sqrt vs sqrt() (development aid; the test tier has the same check)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc
for mode in (20, 21, 22):
    bad, flg = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    rc = lsc.lib().lsc_kat_division(0, 0x5157 + mode, mode, 1 << 15, 256, ctypes.byref(bad), ctypes.byref(flg))
    print(f"sqrt KAT mode {mode}: rc={rc} mismatches={bad.value} flagged={flg.value} of {1 << 31}")
