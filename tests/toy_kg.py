"""Toy knowledge graph for the API tests (plays the role of the reference's tests/data.py:
a dozen string triples + three validation triples, indexed with KGE.data_utils)."""
import numpy as np

from KGE.data_utils import index_kg, convert_kg_to_index

train_raw = np.array([
    ["Curie", "field", "Physics"], ["Curie", "born_in", "Warsaw"], ["Curie", "is_a", "Scientist"],
    ["Noether", "field", "Algebra"], ["Noether", "born_in", "Erlangen"], ["Noether", "is_a", "Scientist"],
    ["Warsaw", "capital_of", "Poland"], ["Erlangen", "located_in", "Germany"], ["Physics", "part_of", "Science"],
    ["Algebra", "part_of", "Mathematics"], ["Hilbert", "colleague_of", "Noether"], ["Hilbert", "is_a", "Scientist"],
    ["Poland", "neighbour_of", "Germany"], ["Mathematics", "related_to", "Science"],
])
val_raw = np.array([["Hilbert", "field", "Algebra"], ["Curie", "is_a", "Scientist"], ["Noether", "born_in", "Erlangen"]])

metadata = index_kg(train_raw)
train = convert_kg_to_index(train_raw, metadata["ent2ind"], metadata["rel2ind"]).astype(np.int64)
val = convert_kg_to_index(val_raw, metadata["ent2ind"], metadata["rel2ind"]).astype(np.int64)
