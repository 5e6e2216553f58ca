"""TEST INFRASTRUCTURE ONLY -- empty stand-in for dm-tree, which the reference imports at
sgnn/single_scale/evaluate.py:3 and never uses."""
