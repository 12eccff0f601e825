"""tesseract_b200 -- B200-native batched collision checking behind the Tesseract contact-manager API."""
