"""gaudi_b200 — B200-native batched genome expression, Contacts scoring and pairwise RMSD
for GaudiMM (insilichem/gaudi), behind the reference's gene/objective/similarity plugin API."""
__version__ = "0.1.0"
