"""Test / bench scaffolding shipped with the package (not part of the decode path): deterministic synthetic inputs,
table-driven full scorers and a small attention decoder / LM with the reference's module surface."""
