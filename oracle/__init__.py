"""Test-infrastructure only: CPU restatement of the reference's hot path (see unet_oracle.py header)."""
