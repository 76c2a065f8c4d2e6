"""Import stub (test infrastructure) for the un-vendored VideoX-Fun submodule the reference imports
(/root/reference/.gitmodules:4-6). Single-process, sequence-parallel world size 1."""
