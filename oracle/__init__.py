"""Test infrastructure only: CPU oracle + bridge to the compiled reference. Never imported by metta_b200."""
