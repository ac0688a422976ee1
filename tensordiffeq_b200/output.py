def print_screen(model, discovery_model=False):
    """tensordiffeq/output.py:5-10 without the pyfiglet banner"""
    print("TensorDiffEq (B200-native engine)")
    if discovery_model:
        print("Running Discovery Model for Parameter Estimation\n\n")
    print("Neural Network Model Summary\n")
    print(model.u_model.summary())
