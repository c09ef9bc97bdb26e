"""b200-ainic: Blackwell-native gradient all-reduce + weight-update engine."""
import os as _os

# The flow all-reduce is a persistent kernel that is launched BEFORE the GEMMs whose output it waits
# for.  With CUDA's default lazy module loading the first launch of a not-yet-loaded kernel has to
# load it while the persistent kernel already occupies the GPU, and that load can wait for the device
# to drain: a dead-lock broken only by the kernel's own time-out.  Eager loading (effective if set
# before the CUDA context exists) removes the hazard; the native trainer additionally launches every
# kernel of a step once at creation time.
_os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
