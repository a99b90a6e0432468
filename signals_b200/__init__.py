"""signals_b200 -- B200 (sm_100a) implementation of the per-cell haplotype- / allele-specific
copy-number HMM of shahcompbio/signals (emission + Viterbi-style decode), behind the reference's
own function names.  See DESIGN.md and INTEGRATION.md.

Importing the package does not load CUDA; the first call into signals_b200.hmm loads
libsignals_b200.so and raises if it has not been built -- there is no CPU fallback.
"""
from . import _lib  # noqa: F401

__all__ = ["hmm", "synth", "_lib"]
__version__ = "0.1.0"
