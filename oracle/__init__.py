"""CPU oracle (test infrastructure). See sysml_oracle.cpp. Never imported by systemml_b200/."""
