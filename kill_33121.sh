ssh test@127.0.0.1 "kill -9 1642"
ssh test@127.0.0.1 "kill -9 1643"
ssh test@127.0.0.1 "kill -9 1644"
kill -9 1591
