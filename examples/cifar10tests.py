#!/usr/bin/env python
"""examples/cifar10tests.lua in the reference is a byte-for-byte copy of examples/cifar10.lua; here the name is kept so
that a command line written for it still works, and it runs the one CIFAR-10 program (examples/cifar10.py)."""
from cifar10 import main

if __name__ == "__main__":
    main()
