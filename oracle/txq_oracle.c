/* placeholder: transform/quant oracle lands here */
