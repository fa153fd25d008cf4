"""CPU oracle for the TP hot path -- TEST INFRASTRUCTURE, never the product path."""
