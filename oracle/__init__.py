"""CPU oracles -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package.  Nothing under spatial_queue_b200/ imports it (tests/test_no_oracle_in_product.py enforces that).

  oracle.sssp   C restatement of sp's Dijkstra (sssp_oracle.c) + loader for the real reference library
                compiled from /root/reference into oracle/_ref/liblsp.so.
  oracle.tick   C restatement of the reference tick (tick_oracle.c).
"""
