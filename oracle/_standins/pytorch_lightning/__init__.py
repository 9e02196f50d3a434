"""Test-infrastructure stand-in: the reference imports pytorch_lightning only for AttributeDict."""
