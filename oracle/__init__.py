"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's algorithm for the modality-embedding hot path.  It exists to check
the CUDA path; it is never the thing shipped or measured as the product.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it.

Parity status: the reference pins nothing for this path (no tests, no golden vectors, transformers
unpinned — SURVEY.md §4, §8 C4).  The oracle is therefore pinned against OUTPUTS OF THE REFERENCE
ITSELF RUN IN THE BUILD CONTAINER: oracle/make_golden.py imports /root/reference (unmodified) plus the
installed transformers 5.5.0, runs ImageModality / MLPProjector / embed_modalities_with_text /
compute_token_range / the collator index build, and commits the results under tests/golden/.
"""
