"""Oracle package -- TEST INFRASTRUCTURE ONLY (see oracle/visitron_oracle.py header).

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs only.  The product package ``visitron_b200`` never imports it.
"""
