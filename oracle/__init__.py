"""CPU oracle for the tachikoma QNN hot path -- TEST INFRASTRUCTURE ONLY (see qnn_oracle.c)."""
