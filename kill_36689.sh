ssh test@127.0.0.1 "kill -9 1972"
ssh test@127.0.0.1 "kill -9 1973"
kill -9 1591
