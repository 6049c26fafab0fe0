// TEST INFRASTRUCTURE — not part of the product (see ref_harness.cc).
// Harness around the reference's ConstrainedNonlinearLeastSquares + transform-chain residuals.
// (filled in with the NLS stage)
