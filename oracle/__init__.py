"""CPU oracle for the cudasp_scan hot path.  TEST INFRASTRUCTURE ONLY (see ref_scan.py header)."""
