"""Test and benchmark harness: everything here may use torch; the product (oni_ml_b200/) does not."""
