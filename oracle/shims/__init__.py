"""Stand-ins for the DGL / PyG *primitives* that generic/eegatconv.py and
generic/gatv3_conv.py call, so those reference files can be executed verbatim in
a container that has neither library (tests/golden/make_golden.py puts this
directory on sys.path).  Only the primitives are restated; the layer code that
runs is the reference's own.  TEST INFRASTRUCTURE."""
