ssh test@127.0.0.1 "kill -9 5096"
ssh test@127.0.0.1 "kill -9 5097"
ssh test@127.0.0.1 "kill -9 5098"
kill -9 5044
