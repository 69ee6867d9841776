from .start import main

raise SystemExit(main())
