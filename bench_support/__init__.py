"""Bench / test support that is NOT part of the product: the deterministic synthetic input generator of
SURVEY 8(d) (uses cv2 on the host).  Imported by bench.py, tests/ and tools/ only."""
