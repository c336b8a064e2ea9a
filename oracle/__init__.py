"""Test oracle — TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's CPU-baseline
legs).  Nothing under the product package imports this."""
