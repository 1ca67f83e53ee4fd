from deepfm_b200.train import main
import sys
sys.exit(main())
