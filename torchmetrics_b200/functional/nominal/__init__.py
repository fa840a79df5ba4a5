"""Nominal association measures on the multiclass confusion-matrix update."""
from torchmetrics_b200.functional.nominal.cramers import cramers_v, cramers_v_matrix
from torchmetrics_b200.functional.nominal.pearson import (
    pearsons_contingency_coefficient,
    pearsons_contingency_coefficient_matrix,
)
from torchmetrics_b200.functional.nominal.theils_u import theils_u, theils_u_matrix
from torchmetrics_b200.functional.nominal.tschuprows import tschuprows_t, tschuprows_t_matrix

__all__ = ["cramers_v", "cramers_v_matrix", "pearsons_contingency_coefficient",
           "pearsons_contingency_coefficient_matrix", "theils_u", "theils_u_matrix", "tschuprows_t",
           "tschuprows_t_matrix"]
