"""TEST INFRASTRUCTURE — parity checkers for the direct-summation path. Never imported by sctl_b200/."""
