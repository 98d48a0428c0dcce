"""oracle — CPU restatement of the hot path.  TEST INFRASTRUCTURE ONLY: imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, never by cekirdekler_b200."""
