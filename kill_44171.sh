ssh test@127.0.0.1 "kill -9 2298"
ssh test@127.0.0.1 "kill -9 2299"
ssh test@127.0.0.1 "kill -9 2300"
kill -9 2109
